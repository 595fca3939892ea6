"""gloo world_size=2 integration tests on CPU (BASELINE config 1): the DDP and FSDP engines must produce the same
parameters as a single process training on the concatenated batch (mean loss over equal-sized rank batches)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, strategy, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from hyperion_b200.models import SimpleTransformerLM
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine

    torch.set_num_threads(1)
    dev = setup(rank, world, "gloo", seed=0)
    torch.manual_seed(0)
    model = SimpleTransformerLM(300, 32, 2, 2, 64, dropout=0.0)
    reshard = strategy == "fsdp_reshard"
    strategy = "fsdp" if reshard else strategy
    eng = build_engine(model, strategy=strategy, precision="fp32", lr=1e-2, clip=1.0 if strategy == "fsdp" else 0.0,
                       device=dev, fsdp_min_num_params=2000, reshard_after_forward=reshard,
                       fsdp_units=list(model.tr.layers) if reshard else None)
    if reshard:
        assert eng.model.reshard and len(eng.model._rs_units) == 2 and eng.model.flat.data.numel() == 0
    g = torch.Generator().manual_seed(1)
    data = torch.randint(0, 300, (3, 2 * world, 12), generator=g)
    losses = []
    for step in range(3):
        b = data[step, rank * 2:(rank + 1) * 2]
        loss = eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous())
        eng.backward_and_step(loss)
        losses.append(float(loss.detach()))
    if reshard:
        # every step gathers each layer for forward; backward finds both layers still in their slots (2 slots, 2 layers)
        assert eng.model.gathers_issued == 3 * 2, eng.model.gathers_issued
    from hyperion_b200.utils.checkpoint import full_state_dict

    sd = full_state_dict(eng.model)
    if rank == 0:
        torch.save({"sd": sd, "losses": losses}, out)
    cleanup()


def _single(strategy):
    from hyperion_b200.models import SimpleTransformerLM

    torch.manual_seed(0)
    model = SimpleTransformerLM(300, 32, 2, 2, 64, dropout=0.0)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
    g = torch.Generator().manual_seed(1)
    data = torch.randint(0, 300, (3, 4, 12), generator=g)
    for step in range(3):
        opt.zero_grad()
        b = data[step]
        # mean of the two rank-local mean losses == what DDP / FSDP average
        loss = 0.5 * (model.loss(b[:2, :-1].contiguous(), b[:2, 1:].contiguous()) + model.loss(b[2:, :-1].contiguous(), b[2:, 1:].contiguous()))
        loss.backward()
        if strategy.startswith("fsdp"):
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
    return model.state_dict()


@pytest.mark.parametrize("strategy", ["ddp", "fsdp", "fsdp_reshard"])
def test_gloo_world2_matches_single_process(tmp_path, strategy):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), strategy, out), nprocs=2, join=True)
    got = torch.load(out, weights_only=False)
    want = _single(strategy)
    assert set(got["sd"]) == set(want)
    for k in want:
        a, b = got["sd"][k], want[k]
        if k.endswith("in_proj_bias"):
            # d(loss)/d(key bias) is analytically zero (softmax is shift-invariant), so Adam normalises pure rounding
            # noise there; compare the query and value thirds only.
            e = a.numel() // 3
            a, b = torch.cat([a[:e], a[2 * e:]]), torch.cat([b[:e], b[2 * e:]])
        assert torch.allclose(a, b, atol=1e-4, rtol=1e-3), k


def _lora_worker(rank, world, port, reshard, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from hyperion_b200.lora import inject_lora
    from hyperion_b200.models import LlamaConfig, LlamaForCausalLM
    from hyperion_b200.models.llama import LlamaDecoderLayer
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine
    from hyperion_b200.utils.checkpoint import full_state_dict

    torch.set_num_threads(1)
    dev = setup(rank, world, "gloo", seed=0)
    torch.manual_seed(0)
    cfg = LlamaConfig.tiny()
    cfg.num_hidden_layers = 3
    model = LlamaForCausalLM(cfg)
    inject_lora(model, r=4, alpha=8, dropout=0.0)
    units = [m for m in model.modules() if isinstance(m, LlamaDecoderLayer)]
    eng = build_engine(model, strategy="fsdp", precision="fp32", lr=1e-2, clip=1.0, device=dev, fsdp_units=units,
                       reshard_after_forward=reshard)
    if reshard:
        f = eng.model
        assert f.reshard and len(f._rs_units) == 3 and f._frozen and f.fshard16 is not None
        # the frozen base of the layers lives in two slots + a 1/W shard, not in 3 full copies
        per_layer = sum(p.numel() for p in units[0].parameters() if not p.requires_grad)
        assert f.fshard16.numel() * world >= 3 * per_layer and f._fslots[0].numel() >= per_layer
    g = torch.Generator().manual_seed(1)
    data = torch.randint(0, cfg.vocab_size, (4, 2 * world, 16), generator=g)
    losses = []
    for step in range(4):
        b = data[step, rank * 2:(rank + 1) * 2]
        loss = eng.model(input_ids=b, labels=b.clone()).loss
        eng.backward_and_step(loss)
        losses.append(float(loss.detach()))
    sd = full_state_dict(eng.model)
    if rank == 0:
        torch.save({"sd": sd, "losses": losses}, out)
    cleanup()


def test_gloo_lora_fsdp_reshard_shards_frozen_base(tmp_path):
    """FULL_SHARD mode (reshard_after_forward) with LoRA: the frozen base weights are sharded and gathered per layer; the
    run must match the params-resident mode step for step and export the same full state_dict."""
    outs = []
    for reshard in (False, True):
        out = str(tmp_path / f"r{int(reshard)}.pt")
        mp.spawn(_lora_worker, args=(2, _free_port(), reshard, out), nprocs=2, join=True)
        outs.append(torch.load(out, weights_only=False))
    a, b = outs
    assert a["losses"][-1] < a["losses"][0]
    assert torch.allclose(torch.tensor(a["losses"]), torch.tensor(b["losses"]), atol=1e-5), (a["losses"], b["losses"])
    assert set(a["sd"]) == set(b["sd"])
    for k in a["sd"]:
        assert torch.allclose(a["sd"][k], b["sd"][k], atol=1e-5), k


def _resume_worker(rank, world, port, base, phase, strategy):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import glob

    from hyperion_b200 import train

    torch.set_num_threads(1)
    kw = dict(backend="gloo", precision="fp32", synthetic=True, rows=32, batch_size=4, seed=0, save=True,
              model_kwargs=dict(emb_dim=32, n_heads=2, n_layers=1, dim_feedforward=64, dropout=0.0), vocab_size=300, pad_id=299)
    fn = train.train_language_model_ddp if strategy == "ddp" else train.train_language_model_fsdp
    if strategy == "fsdp":
        kw["min_num_params"] = 2000
    if phase == "full":
        hist = fn(rank, world, epochs=3, base_dir=os.path.join(base, "full"), **kw)
    elif phase == "first":
        hist = fn(rank, world, epochs=2, base_dir=os.path.join(base, "part"), **kw)
    else:
        state = sorted(glob.glob(os.path.join(base, "part", "data/distributed/*_train_state.pt")))[-1]
        hist = fn(rank, world, epochs=3, base_dir=os.path.join(base, "part"), resume=state, **kw)
    if rank == 0:
        torch.save(hist, os.path.join(base, f"hist_{phase}.pt"))


@pytest.mark.parametrize("strategy", ["ddp", "fsdp"])
def test_gloo_kill_and_resume_reproduces_the_loss_curve(tmp_path, strategy):
    """Three epochs in one go == two epochs, process exit, ``resume=<train_state.pt>`` in fresh processes for the third
    (weights, fp32 master, optimizer moments -- sharded per rank under FSDP -- and epoch are restored)."""
    base = str(tmp_path)
    for phase in ("full", "first", "resume"):
        mp.spawn(_resume_worker, args=(2, _free_port(), base, phase, strategy), nprocs=2, join=True)
    full = torch.load(os.path.join(base, "hist_full.pt"))
    first = torch.load(os.path.join(base, "hist_first.pt"))
    resumed = torch.load(os.path.join(base, "hist_resume.pt"))
    assert len(full) == 3 and len(first) == 2 and len(resumed) == 1
    assert abs(full[2] - resumed[0]) < 1e-4 * max(1.0, abs(full[2])), (full, first, resumed)


def _reshard_ckpt_worker(rank, world, port, reshard, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from hyperion_b200.models import GPT2Config, GPT2LMHeadModel
    from hyperion_b200.models.gpt2 import Block
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine
    from hyperion_b200.utils.checkpoint import full_state_dict

    torch.set_num_threads(1)
    dev = setup(rank, world, "gloo", seed=0)
    torch.manual_seed(0)
    model = GPT2LMHeadModel(GPT2Config(vocab_size=256, n_positions=32, n_embd=32, n_layer=3, n_head=2))
    model.use_checkpoint = True            # every block is recomputed inside backward: gathers fire from the recompute too
    eng = build_engine(model, strategy="fsdp", precision="fp32", lr=1e-2, clip=1.0, device=dev,
                       fsdp_units=[m for m in model.modules() if isinstance(m, Block)], reshard_after_forward=reshard)
    eng.model.train()
    g = torch.Generator().manual_seed(3)
    data = torch.randint(0, 256, (4, 2 * world, 24), generator=g)
    losses = []
    for step in range(4):
        if step == 2:
            # weights-only reload in the middle of training: shards / master (and, when resharded, the slots) must follow
            sd = full_state_dict(eng.model)
            eng.model.load_state_dict(sd)
        loss = eng.model.loss(data[step, rank * 2:(rank + 1) * 2])
        eng.backward_and_step(loss)
        losses.append(float(loss.detach()))
    sd = full_state_dict(eng.model)
    if rank == 0:
        torch.save({"sd": sd, "losses": losses}, out)
    cleanup()


def test_gloo_full_shard_with_activation_checkpointing_and_reload(tmp_path):
    """FULL_SHARD slots + activation checkpointing (block forwards re-run inside backward) + a mid-run load_state_dict:
    same losses and parameters as the params-resident mode."""
    outs = []
    for reshard in (False, True):
        out = str(tmp_path / f"c{int(reshard)}.pt")
        mp.spawn(_reshard_ckpt_worker, args=(2, _free_port(), reshard, out), nprocs=2, join=True)
        outs.append(torch.load(out, weights_only=False))
    a, b = outs
    assert torch.allclose(torch.tensor(a["losses"]), torch.tensor(b["losses"]), atol=1e-5), (a["losses"], b["losses"])
    for k in a["sd"]:
        assert torch.allclose(a["sd"][k], b["sd"][k], atol=1e-5), k


class _PartlyUsed(torch.nn.Module):
    """Two branches; only `a` is used in forward, so `b` never receives a gradient."""

    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(16, 16)
        self.b = torch.nn.Linear(16, 16)

    def forward(self, x):
        return self.a(x)


def _unused_worker(rank, world, port, strategy, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine

    torch.set_num_threads(1)
    dev = setup(rank, world, "gloo", seed=0)
    torch.manual_seed(0)
    model = _PartlyUsed()
    # default 25 MB buckets / one root unit: the used and the unused branch share a bucket (unit), whose hook counter
    # therefore never reaches zero on its own
    eng = build_engine(model, strategy=strategy, precision="fp32", lr=1e-2, device=dev, fsdp_min_num_params=10 ** 6)
    g = torch.Generator().manual_seed(10 + rank)          # different data per rank: only the all-reduce keeps them equal
    for _ in range(3):
        x = torch.randn(4, 16, generator=g)
        eng.backward_and_step(eng.model(x).pow(2).mean())
    from hyperion_b200.utils.checkpoint import full_state_dict

    sd = full_state_dict(eng.model)
    flat = torch.cat([v.reshape(-1) for v in sd.values()])
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    torch.distributed.all_gather(gathered, flat)
    if rank == 0:
        torch.save({"same": bool(torch.allclose(gathered[0], gathered[1], atol=1e-6)),
                    "unused": list(getattr(eng.model, "unused", [])), "moved": float((sd["a.weight"] - _PartlyUsed().a.weight).abs().sum())}, out)
    cleanup()


@pytest.mark.parametrize("strategy", ["ddp", "fsdp"])
def test_gloo_parameters_without_gradient_do_not_desynchronise_the_ranks(tmp_path, strategy):
    """A bucket / unit holding a parameter that got no gradient is still reduced at the end of backward (with zeros in the
    missing slots): the ranks stay identical, the used branch trains, and the engine names the unused parameters."""
    out = str(tmp_path / "u.pt")
    mp.spawn(_unused_worker, args=(2, _free_port(), strategy, out), nprocs=2, join=True)
    r = torch.load(out, weights_only=False)
    assert r["same"], "ranks diverged"
    assert any(n.startswith("b.") for n in r["unused"]), r["unused"]


def test_launch_script_runs_the_fsdp_job_under_torchrun(tmp_path):
    """scripts/run_language_fsdp.sh (reference: 02_development/run_language_fsdp.sh) end to end on gloo, 2 ranks: the launcher, the
    environment preset, run_distributed.py, the FSDP trainer, the metrics CSV and the checkpoint."""
    import subprocess
    from pathlib import Path

    import pandas as pd

    root = Path(__file__).resolve().parent.parent
    env = {**os.environ, "NGPU": "2", "EPOCHS": "2", "BASE_DIR": str(tmp_path), "CUDA_VISIBLE_DEVICES": ""}
    r = subprocess.run(["bash", str(root / "scripts/run_language_fsdp.sh"), "--backend", "gloo", "--steps", "2", "--batch_size", "4"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    csvs = [f for f in (tmp_path / "data/distributed").glob("language_fsdp_2gpus_*_metrics.csv") if "device_metrics" not in f.name]
    assert len(csvs) == 1, list((tmp_path / "data/distributed").iterdir())
    df = pd.read_csv(csvs[0])
    assert list(df["epoch"]) == [1, 2] and set(df["gpus"]) == {2} and {"loss", "duration"} <= set(df.columns)
    assert any((tmp_path / "data/distributed").glob("*.pt"))


def _eval_between_worker(rank, world, port, strategy, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from hyperion_b200.models import SimpleTransformerLM
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine

    torch.set_num_threads(1)
    dev = setup(rank, world, "gloo", seed=0)
    torch.manual_seed(0)
    model = SimpleTransformerLM(300, 32, 2, 2, 64, dropout=0.0)
    reshard = strategy == "fsdp_reshard"
    eng = build_engine(model, strategy="fsdp" if reshard else strategy, precision="fp32", lr=1e-2, device=dev,
                       fsdp_min_num_params=2000, reshard_after_forward=reshard, fsdp_units=list(model.tr.layers) if reshard else None)
    data = torch.randint(0, 300, (4, 2 * world, 12), generator=torch.Generator().manual_seed(1))
    val = []
    for step in (0, 1, 2, 0):
        b = data[step, rank * 2:(rank + 1) * 2]
        eng.model.train()
        if step == 0:
            eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous())      # a forward whose graph is dropped without a backward
        eng.backward_and_step(eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous()))
        eng.model.eval()                                                       # validation between steps: no_grad, twice
        with torch.no_grad():
            v = data[3, :2]
            l1 = float(eng.model.loss(v[:, :-1].contiguous(), v[:, 1:].contiguous()))
            l2 = float(eng.model.loss(v[:, :-1].contiguous(), v[:, 1:].contiguous()))
        assert l1 == l2
        val.append(l1)
    if rank == 0:
        torch.save(val, out)
    cleanup()


@pytest.mark.parametrize("strategy", ["ddp", "fsdp", "fsdp_reshard"])
def test_gloo_validation_passes_between_steps(tmp_path, strategy):
    """eval()/no_grad forwards between optimizer steps (a validation loop) and a forward that is never backpropagated leave the engines
    consistent: the validation loss after every step equals the single-process run's."""
    from hyperion_b200.models import SimpleTransformerLM

    torch.manual_seed(0)
    model = SimpleTransformerLM(300, 32, 2, 2, 64, dropout=0.0)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
    data = torch.randint(0, 300, (4, 4, 12), generator=torch.Generator().manual_seed(1))
    want = []
    for step in (0, 1, 2, 0):
        model.train()
        b = data[step]
        opt.zero_grad()
        model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous()).backward()
        opt.step()
        model.eval()
        with torch.no_grad():
            v = data[3, :2]
            want.append(float(model.loss(v[:, :-1].contiguous(), v[:, 1:].contiguous())))
    out = tmp_path / "val.pt"
    mp.spawn(_eval_between_worker, args=(2, _free_port(), strategy, str(out)), nprocs=2, join=True)
    got = torch.load(out)
    assert got == pytest.approx(want, rel=2e-5), (got, want)


def test_model_all_runs_the_four_jobs_in_one_process_group(tmp_path):
    """`run_distributed.py --model all` (reference: run_distributed.py:134-139): four trainers back to back in one torchrun launch.  The
    process group is held across them (tearing it down and re-creating it in the same process fails on gloo)."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    cmd = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", "2",
           str(root / "run_distributed.py"), "--model", "all", "--llama-size", "tiny", "--epochs", "1", "--steps", "2", "--batch_size", "2",
           "--backend", "gloo", "--base_dir", str(tmp_path), "--no-save"]
    r = subprocess.run(cmd, env={**os.environ, "CUDA_VISIBLE_DEVICES": ""}, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    for tag in ("[LM-DDP] Epoch 1/1", "[CIFAR-DDP] Epoch 1/1", "[LM-FSDP] Epoch 1/1", "[Llama-fsdp_bf16] Epoch 1/1"):
        assert tag in r.stdout, tag
    kinds = sorted({f.name.split("_2gpus_")[0] for f in (tmp_path / "data/distributed").glob("*_metrics.csv")})
    assert kinds == ["cifar_ddp", "language_ddp", "language_fsdp", "llama"]
    assert (tmp_path / "data/distributed/scaling_analysis.csv").exists()


def test_hold_and_release_of_the_process_group():
    from hyperion_b200.parallel import runtime

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()), RANK="0", WORLD_SIZE="1")
    runtime.hold_process_group()
    try:
        runtime.setup(0, 1, "gloo")
        runtime.cleanup()
        assert torch.distributed.is_initialized()          # held: cleanup() only synchronised
        runtime.setup(0, 1, "gloo")                        # a second trainer's setup() reuses the group
        runtime.cleanup()
        assert torch.distributed.is_initialized()
    finally:
        runtime.release_process_group()
    assert not torch.distributed.is_initialized()
