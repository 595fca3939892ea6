"""Multi-GPU parity check (run under torchrun): trains the same model for a few steps with the fused NVLink
communication kernels and with the NCCL baseline path, from identical weights and data, and compares the loss curves and
final parameters.  ``python -m torch.distributed.run --nproc-per-node 2 -m hyperion_b200.parallel.parity --strategy ddp``.
Exit code 0 = parity within tolerance on every rank.  ("DDP/FSDP loss-curve parity fused-vs-baseline".)"""
from __future__ import annotations

import argparse
import copy
import os
import sys

import torch
import torch.distributed as dist


def run(strategy: str, precision: str, steps: int, reshard: bool = False, model_name: str = "lm") -> int:
    from hyperion_b200.models import SimpleTransformerLM
    from hyperion_b200.parallel.runtime import cleanup, setup
    from hyperion_b200.train.engine import build_engine
    from hyperion_b200.utils.checkpoint import full_state_dict

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = setup(rank, world, "nccl", seed=0)
    torch.manual_seed(0)
    gpt2 = model_name == "gpt2"
    if gpt2:
        os.environ.setdefault("HYPERION_FSDP_AG_BLOCKS", "1")   # exercise the gather-inside-GEMM path on block units
        # block-sized units: every Block is gathered inside its c_attn GEMM (K-C3 on composite units), the root unit (wte,
        # wpe, ln_f and the carved-out block norms) by the stand-alone kernel -- the mix the BASELINE GPT-2 config runs
        from hyperion_b200.models import GPT2Config, GPT2LMHeadModel
        from hyperion_b200.models.gpt2 import Block

        base = GPT2LMHeadModel(GPT2Config(vocab_size=1024, n_positions=128, n_embd=256, n_layer=3, n_head=4))
        data = torch.randint(0, 1024, (steps, world, 4, 128), generator=torch.Generator().manual_seed(7)).to(dev)
    else:
        base = SimpleTransformerLM(50257, dropout=0.0)
        data = torch.randint(0, 50256, (steps, world, 8, 64), generator=torch.Generator().manual_seed(7)).to(dev)
    results = {}
    for fused in (True, False):
        model = copy.deepcopy(base)
        # --reshard: the fused arm runs FULL_SHARD (encoder layers as units, gathered into two slots before forward and again
        # before backward); the NCCL arm keeps the parameters resident -- the two must still agree
        rs = reshard and fused
        eng = build_engine(model, strategy=strategy, precision=precision, lr=1e-3, clip=1.0 if strategy == "fsdp" else 0.0,
                           device=dev, fused_comm=fused, reshard_after_forward=rs,
                           fsdp_units=([m for m in model.modules() if isinstance(m, Block)] if gpt2 else
                                       list(model.tr.layers) if (reshard and strategy == "fsdp") else None))
        assert getattr(eng.model, "fused", None) == fused, "requested comm path was not taken"
        if strategy == "fsdp" and fused and not reshard:
            assert len(eng.model._ag_handles) > 0, "no unit uses the all-gather-fused GEMM"
        if rs:
            assert eng.model.reshard and len(eng.model._rs_units) == (3 if gpt2 else 2) and eng.model.flat.data.numel() == 0
        if gpt2 and fused and not reshard and strategy == "fsdp":
            assert len(eng.model._ag_handles) == 3, "the three Block units must gather inside their first GEMM"
        losses = []
        for s in range(steps):
            b = data[s, rank]
            loss = (eng.model.loss(b) if gpt2 else
                    eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous(), ignore_index=50256))
            eng.backward_and_step(loss)
            losses.append(loss.detach().float())
        torch.cuda.synchronize()
        results[fused] = (torch.stack(losses).cpu(), full_state_dict(eng.model))
        del eng, model
    lf, sf = results[True]
    ln, sn = results[False]
    ok = bool(torch.isfinite(lf).all() and torch.isfinite(ln).all())   # NaN = a poisoned (not yet gathered) buffer was read
    ldiff = (lf - ln).abs().max().item()
    if ldiff > 5e-2:
        ok = False
    worst, worst_key = 0.0, ""
    for k in sn:
        if k.endswith("in_proj_bias"):
            continue  # the key-bias gradient is analytically zero; Adam turns rounding noise into +-lr steps there
        a, b = sf[k].float(), sn[k].float()
        if k.endswith("c_attn.bias"):   # same for the key third of GPT-2's packed qkv bias
            e = a.numel() // 3
            a, b = torch.cat([a[:e], a[2 * e:]]), torch.cat([b[:e], b[2 * e:]])
        # denominator: the tensor's norm plus the distance Adam can move it at all in `steps` steps of lr 1e-3 -- zero-initialised
        # biases are pure +-lr sign steps of near-zero gradients, where the summation order of the 16-bit reduction decides
        rel = ((a - b).norm() / (b.norm() + steps * 1e-3 * b.numel() ** 0.5)).item()
        if a.dim() <= 1:
            rel *= 0.4    # vectors (biases, norm scale/shift) are allowed 5e-2 instead of 2e-2: few elements, mostly sign steps
        if rel > worst:
            worst, worst_key = rel, k
    if worst > 2e-2:
        ok = False
    print(f"[rank {rank}] {strategy}/{precision}: fused losses {lf.tolist()} nccl losses {ln.tolist()} max|dloss|={ldiff:.4f} "
          f"worst param rel diff={worst:.2e} ({worst_key}) -> {'PASS' if ok else 'FAIL'}", flush=True)
    bad = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(bad)
    cleanup()
    return int(bad.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--strategy", default="ddp", choices=["ddp", "fsdp"])
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--reshard", action="store_true")
    ap.add_argument("--model", default="lm", choices=["lm", "gpt2"])
    a = ap.parse_args()
    sys.exit(1 if run(a.strategy, a.precision, a.steps, a.reshard, a.model) else 0)


if __name__ == "__main__":
    main()
