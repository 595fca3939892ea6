"""``bench.py --impl reference``: drives the UNMODIFIED reference trainers installed in ``baseline/_ref``.

The reference (``/root/reference/02_development/distributed_utils.py``) is a module of trainer functions that read an
on-disk dataset, build their own model / wrapper / optimizer and loop over a ``DataLoader``.  This file changes none of
that.  It only supplies the environment those functions expect, and measures them from the outside:

* import stand-ins for ``matplotlib`` and ``peft``: the module imports both at the top (distributed_utils.py:26, 35) but
  the language / CIFAR trainers never touch them; neither package is in this image;
* a working directory holding ``gpt2/tokenizer.json`` so that ``AutoTokenizer.from_pretrained("gpt2")``
  (distributed_utils.py:154, 311) resolves offline to a 50257-entry vocabulary with eos = pad = 50256;
* ``<base_dir>/data/processed/wikitext2_tokenized`` / ``cifar10_train.pt`` in the reference's formats
  (distributed_utils.py:149-152, 222), synthetic, sized for exactly the warm-up + timed steps asked for;
* timing: a process-wide ``nn.Module`` forward pre-hook (public torch API) counts top-level forward calls = loop
  iterations.  CUDA events are recorded at the start of iteration W and of iteration W+K, so the timed region is K
  complete iterations of the reference's own loop -- DataLoader fetch, host->device copy, forward, backward, collective,
  optimizer step and ``loss.item()``.  After the second event the hook raises a private exception that unwinds out of the
  trainer (the reference's end-of-run checkpoint is not part of the metric).  The number is therefore END-TO-END by
  construction; it is reported both as ``value`` and as ``e2e.value`` and flagged ``"timed": "end_to_end"``.

Jobs available:  lm_ddp -> train_language_model_ddp (distributed_utils.py:132-200), lm_fsdp -> train_language_model_fsdp
(:290-406), cifar_ddp -> train_cifar_model_ddp (:207-279).  ``train_llama_fsdp`` (:413-545; the only path that could load
GPT-2 or Llama) calls ``from_pretrained(..., device_map=...)``, which transformers refuses without ``accelerate``, and its
LoRA branch needs ``peft``; neither is installed and there is no index, so gpt2_fsdp / llama_lora print ``unavailable``
and name ``--impl torch`` as the operative baseline.  The reference contains no ResNet-50 / ViT distributed job at all.
"""
from __future__ import annotations

import json
import os
import sys
import time
import types
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"

JOBS = {
    "lm_ddp": ("train_language_model_ddp", "SimpleTransformerLM DDP tokens/s", "fp16"),
    "lm_fsdp": ("train_language_model_fsdp", "SimpleTransformerLM FSDP tokens/s", "bf16"),
    "cifar_ddp": ("train_cifar_model_ddp", "ResNet-18 CIFAR DDP samples/s", "fp16"),
}
NOT_IN_REFERENCE = {
    "gpt2_fsdp": "the reference loads GPT-2/Llama only through train_llama_fsdp (distributed_utils.py:413), whose "
                 "from_pretrained(device_map=...) needs `accelerate` (absent, no index)",
    "llama_lora": "train_llama_fsdp(lora=True) needs `peft` and `accelerate` (both absent, no index)",
    "llama_fsdp": "train_llama_fsdp needs `accelerate` for from_pretrained(device_map=...) (absent, no index)",
    "resnet50_ddp": "the reference has no ResNet-50 distributed job (its DDP vision job is ResNet-18/CIFAR-10: --config cifar_ddp)",
    "vit_ddp": "the reference has no ViT distributed job",
}


class _BenchDone(Exception):
    pass


def _unavailable(why: str, config: str) -> int:
    if int(os.environ.get("RANK", 0)) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why, "config": {"name": config}, "same_config": False,
                          "operative_baseline": "bench.py --impl torch --config " + config,
                          "reference_configs": sorted(JOBS),
                          "how_to_measure_the_reference": "python bench.py --impl reference --config lm_ddp|lm_fsdp|cifar_ddp "
                                                          "(same flags for the product arm); committed runs of both arms at "
                                                          "N=1/2/4/8: profiles/bench/r2/matrix_n<N>_<config>_{reference,hyperion}.json, "
                                                          "table in BASELINE.md section 3"}), flush=True)
    return 0


def _shim_modules() -> list[str]:
    """Stand-ins for modules the reference imports at module scope but does not use on the paths driven here."""
    made = []
    for name in ("matplotlib", "matplotlib.pyplot", "peft"):
        try:
            __import__(name)
        except Exception:  # noqa: BLE001
            m = types.ModuleType(name)
            if name == "peft":
                for attr in ("LoraConfig", "get_peft_model", "TaskType"):
                    def _absent(*a, _n=attr, **k):
                        raise RuntimeError(f"peft.{_n} is a stand-in: peft is not installed in this image")
                    setattr(m, attr, _absent)
            sys.modules[name] = m
            if "." in name:
                setattr(sys.modules[name.split(".")[0]], name.split(".")[1], m)
            made.append(name)
    return made


def _write_tokenizer(dst: Path) -> None:
    from tokenizers import Tokenizer
    from tokenizers.models import WordLevel

    dst.mkdir(parents=True, exist_ok=True)
    vocab = {f"t{i}": i for i in range(50256)}
    vocab["<|endoftext|>"] = 50256
    Tokenizer(WordLevel(vocab, unk_token="<|endoftext|>")).save(str(dst / "tokenizer.json"))
    (dst / "tokenizer_config.json").write_text(json.dumps({
        "tokenizer_class": "PreTrainedTokenizerFast", "eos_token": "<|endoftext|>", "bos_token": "<|endoftext|>",
        "unk_token": "<|endoftext|>", "model_max_length": 1024}))


def _prepare(work: Path, config: str, rows: int, seq: int) -> None:
    """Rank 0 writes tokenizer + dataset; the other ranks wait for the marker file (the process group does not exist yet:
    the reference's own ``setup()`` creates it)."""
    sys.path.insert(0, str(HERE.parent))
    from hyperion_b200.data import synth

    marker = work / ".ready"
    if int(os.environ.get("RANK", 0)) == 0:
        _write_tokenizer(work / "gpt2")
        proc = work / "data" / "processed"
        proc.mkdir(parents=True, exist_ok=True)
        if config == "cifar_ddp":
            import torch

            torch.save(synth.cifar10_like_list(rows, seed=0), proc / "cifar10_train.pt")
        else:
            synth.write_wikitext2_like(proc / "wikitext2_tokenized", seq_len=seq, rows={"train": rows})
        marker.write_text("ok")
    else:
        t0 = time.time()
        while not marker.exists():
            if time.time() - t0 > 600:
                raise TimeoutError("reference arm: rank 0 did not finish preparing the dataset")
            time.sleep(0.2)


def run(args) -> int:
    config = args.config
    if config not in JOBS:
        return _unavailable(NOT_IN_REFERENCE.get(config, f"no reference job for {config}"), config)
    if not (REF_DIR / "distributed_utils.py").exists():
        return _unavailable("baseline/_ref is empty: run baseline/install_reference.sh (needs /root/reference)", config)

    import torch

    if not torch.cuda.is_available():
        return _unavailable("no CUDA device (the reference trainers call torch.cuda.set_device)", config)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29517")
    os.environ["HF_HUB_OFFLINE"] = "1"
    os.environ["TRANSFORMERS_OFFLINE"] = "1"
    os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")
    fn_name, metric, dtype = JOBS[config]
    W, K = args.warmup, args.steps
    # the reference hard-codes its per-GPU batch (32 for the LMs, 64 for CIFAR) and reads seq 128 rows
    bs = 64 if config == "cifar_ddp" else 32
    seq = 128
    rows = (W + K + 2) * bs * world
    work = Path(f"/tmp/hyperion_refarm_{os.environ['MASTER_PORT']}_{config}_{world}_{W}_{K}")
    if rank == 0 and work.exists():
        import shutil

        shutil.rmtree(work)
    if rank != 0:
        time.sleep(1.0)   # let rank 0 clear a stale directory before looking for the marker
    work.mkdir(parents=True, exist_ok=True)
    _prepare(work, config, rows, seq)
    shims = _shim_modules()
    os.chdir(work)                     # "gpt2" resolves to ./gpt2
    sys.path.insert(0, str(REF_DIR))
    import distributed_utils as ref    # the unmodified reference module

    assert Path(ref.__file__).resolve().parent == REF_DIR.resolve(), ref.__file__

    from hyperion_b200.utils.timing import ClockSampler   # measurement only

    st = {"depth": 0, "iters": 0, "e0": None, "e1": None, "h0": 0.0, "sampler": None, "ms": None, "host_ms": None}

    def pre_hook(module, inputs):
        st["depth"] += 1
        if st["depth"] != 1:
            return
        st["iters"] += 1
        if st["iters"] == W + 1:
            torch.cuda.synchronize()
            if rank == 0:
                st["sampler"] = ClockSampler(torch.cuda.current_device()).start()
            st["e0"] = torch.cuda.Event(enable_timing=True)
            st["e0"].record()
            st["h0"] = time.perf_counter()
        elif st["iters"] == W + K + 1:
            st["e1"] = torch.cuda.Event(enable_timing=True)
            st["e1"].record()
            torch.cuda.synchronize()
            st["host_ms"] = (time.perf_counter() - st["h0"]) * 1e3
            st["ms"] = st["e0"].elapsed_time(st["e1"])
            raise _BenchDone()

    def post_hook(module, inputs, output):
        st["depth"] -= 1

    h1 = torch.nn.modules.module.register_module_forward_pre_hook(pre_hook)
    h2 = torch.nn.modules.module.register_module_forward_hook(post_hook, always_call=True)
    try:
        getattr(ref, fn_name)(rank, world, epochs=1, base_dir=str(work))
        done = False
    except _BenchDone:
        done = True
    finally:
        h1.remove()
        h2.remove()
    import torch.distributed as dist

    if not done:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"{fn_name} returned after {st['iters']} iterations, "
                              f"before the {W}+{K} asked for"}), flush=True)
        return 0
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.tensor([st["ms"], st["host_ms"]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0].item()) / K
    clocks = st["sampler"].stop() if st["sampler"] is not None else None
    items = bs * (seq - 1) if config != "cifar_ddp" else bs
    unit = "samples/s" if config == "cifar_ddp" else "tokens/s"
    h2d = bs * seq * 8 if config != "cifar_ddp" else bs * (3 * 32 * 32 * 4 + 8)
    d2h = 4 if config != "cifar_ddp" else 12   # loss.item() (+ correct-count .item() in the CIFAR loop)
    value = items * world / (ms / 1e3)
    if rank == 0:
        cfg = ({"model": "SimpleTransformerLM(50257,256,4,2)", "global_batch": bs * world, "seq_len": seq - 1,
                "parallelism": ("dp" if config == "lm_ddp" else "fsdp") + str(world)} if config != "cifar_ddp" else
               {"model": "resnet18(num_classes=10)", "global_batch": bs * world, "image": 32, "parallelism": f"dp{world}"})
        cfg.update({"entry": f"distributed_utils.{fn_name}", "dataloader": "reference's own (num_workers=2, pageable memory)",
                    "import_stand_ins": shims})
        print(json.dumps({"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": K, "warmup": W,
                          "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
                          "data": "synthetic", "impl": "reference", "timed": "end_to_end", "config": cfg, "clocks": clocks,
                          "e2e": {"value": value, "unit": unit, "ms_per_step": ms, "h2d_bytes_per_step": h2d,
                                  "d2h_bytes_per_step": d2h},
                          "host_ms_per_step": float(t[1].item()) / K, "gpu_launches": 0,
                          "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 2 ** 30}), flush=True)
    try:
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass
    return 0
