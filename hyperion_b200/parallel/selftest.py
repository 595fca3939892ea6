"""Multi-GPU self-test of the communication kernels (run under torchrun, one rank per GPU):

    torchrun --standalone --nproc-per-node 2 -m hyperion_b200.parallel.selftest [--bench]

Checks, against NCCL on the same buffers: the fused all-reduce (K-C1; several sizes incl. ragged
ones, fp32/bf16/fp16, AMP unscale, found_inf propagation, back-to-back buckets), the reduce-scatter
with sum-of-squares and with the fused AdamW update (K-C2), and the all-gather (K-C3).  This also is
the ``hyperion comm-check`` command (the reference's test_nccl.py only prints an all-reduce result;
02_development/test_nccl.py:24-44).  Exit code 0 = all checks passed on every rank.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def run(bench: bool = False, out_json: str | None = None) -> int:
    from hyperion_b200.parallel.runtime import setup, cleanup
    from hyperion_b200.parallel.symm import SymmetricArena

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if not torch.cuda.is_available():
        if rank == 0:
            print("comm-check tests the peer-memory communication kernels and needs one CUDA GPU per rank; on a machine without GPUs "
                  "`python test_nccl.py` under torchrun checks the process group over gloo", file=sys.stderr)
        return 2
    dev = setup(rank, world, "nccl")
    torch.manual_seed(1234 + rank)
    arena = SymmetricArena(dev)
    N = 64 * 1024 * 1024 if bench else 4 * 1024 * 1024
    bufs = {dt: arena.alloc(N, dt, dev) for dt in (torch.float32, torch.bfloat16, torch.float16)}
    shard16 = arena.alloc(N // world, torch.bfloat16, dev)
    arena.rendezvous()
    fails = []
    info = {"rank": rank, "world": world, "multicast": arena.has_multicast,
            "provider": "ipc" if arena._ipc_allocs else "torch_symm"}

    def check(name, ok, detail=""):
        if not ok:
            fails.append(f"{name}: {detail}")

    # sanity of NCCL itself (the reference's test_nccl.py, but asserted)
    t = torch.ones(1, device=dev) * rank
    dist.all_reduce(t)
    check("nccl_allreduce", t.item() == world * (world - 1) / 2, t.item())

    # ---- K-C1 all-reduce vs NCCL
    for dt, tol in ((torch.float32, 1e-6), (torch.bfloat16, 1e-2), (torch.float16, 2e-3)):
        buf = bufs[dt]
        for (start, n) in ((0, 1024), (4096, 1000003 // 8 * 8 + 8), (64, N - 64), (0, 8), (128, 520)):
            n = min(n, N - start)
            src = torch.randn(n, device=dev)
            buf[start:start + n].copy_(src.to(dt))
            ref = buf[start:start + n].float().clone()
            dist.all_reduce(ref)
            ref /= world
            arena.found_inf.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            arena.allreduce_avg_(buf, start, start + n)
            torch.cuda.synchronize()
            check(f"allreduce[{dt},{n}]", _rel(buf[start:start + n], ref) < tol, _rel(buf[start:start + n], ref))
            check(f"allreduce_noinf[{dt},{n}]", arena.found_inf.item() == 0.0)
    # back-to-back buckets + AMP unscale + inf propagation
    buf = bufs[torch.float16]
    src = torch.randn(3, 65536, device=dev)
    buf[:3 * 65536].copy_((src * 128.0).reshape(-1).to(torch.float16))
    ref = buf[:3 * 65536].float().clone()
    dist.all_reduce(ref)
    ref = ref / world / 128.0
    inv = torch.tensor([1.0 / 128.0], device=dev)
    dist.barrier()
    for k in range(3):
        arena.allreduce_avg_(buf, k * 65536, (k + 1) * 65536, inv_scale=inv)
    torch.cuda.synchronize()
    check("allreduce_back_to_back_unscale", _rel(buf[:3 * 65536], ref) < 2e-3, _rel(buf[:3 * 65536], ref))
    buf[:4096].copy_(torch.randn(4096, device=dev).to(torch.float16))
    if rank == world - 1:
        buf[777] = float("inf")
    dist.barrier()
    arena.allreduce_avg_(buf, 0, 4096)
    torch.cuda.synchronize()
    dist.barrier()
    check("found_inf_everywhere", arena.found_inf.item() == 1.0, arena.found_inf.item())
    arena.found_inf.zero_()

    # ---- K-C2 reduce-scatter (+ sumsq) vs NCCL, then fused AdamW vs torch AdamW on the shard
    g16 = bufs[torch.bfloat16]
    per = (N // world) // 8 * 8
    g16[:per * world].copy_(torch.randn(per * world, device=dev).to(torch.bfloat16))
    full = g16[:per * world].float().clone()
    dist.all_reduce(full)
    full /= world
    ref_shard = full[rank * per:(rank + 1) * per]
    gshard = torch.empty(per, device=dev)
    ss = torch.zeros(2, device=dev)
    dist.barrier()
    arena.reduce_scatter(g16, 0, per, gshard=gshard, sumsq=ss)
    torch.cuda.synchronize()
    check("reduce_scatter", _rel(gshard, ref_shard) < 1e-2, _rel(gshard, ref_shard))
    check("reduce_scatter_sumsq", abs(ss[0].item() - (gshard.double() ** 2).sum().item()) / ss[0].item() < 1e-3)
    master = torch.randn(per, device=dev)
    m, v = torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    refp = torch.nn.Parameter(master.clone())
    opt = torch.optim.AdamW([refp], lr=1e-3, weight_decay=0.01)
    refp.grad = gshard.clone()
    opt.step()
    dist.barrier()
    arena.reduce_scatter(g16, 0, per, adam=dict(master=master, m=m, v=v, p16=shard16[:per], lr=1e-3, betas=(0.9, 0.999),
                                                eps=1e-8, weight_decay=0.01, step=1))
    torch.cuda.synchronize()
    check("rs_adamw_master", _rel(master, refp.data) < 1e-4, _rel(master, refp.data))
    check("rs_adamw_p16", torch.equal(shard16[:per], master.to(torch.bfloat16)))

    # ---- K-C3 all-gather of the bf16 shards
    dst = torch.empty(per * world, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(world, device=dev, dtype=torch.int32)
    ref_list = [torch.empty(per, device=dev, dtype=torch.bfloat16) for _ in range(world)]
    dist.all_gather(ref_list, shard16[:per].contiguous())
    dist.barrier()
    arena.allgather(shard16, 0, per, dst, chunk_flags=flags)
    torch.cuda.synchronize()
    check("allgather", torch.equal(dst, torch.cat(ref_list)))
    check("allgather_flags", bool((flags > 0).all()))

    # ---- K-C3 fused: all-gather inside the consumer GEMM vs (NCCL all-gather, then matmul)
    torch.manual_seed(99)  # identical on every rank
    Nw, Kw, Mx = 1536, 1024, 2048
    unit_elems = ((Nw * Kw + Nw + world * 128 - 1) // (world * 128)) * (world * 128)
    per_u = unit_elems // world
    full_ref = torch.zeros(unit_elems, device=dev, dtype=torch.bfloat16)
    full_ref[:Nw * Kw + Nw].copy_((torch.randn(Nw * Kw + Nw, device=dev) * 0.05).to(torch.bfloat16))
    shard16[:per_u].copy_(full_ref[rank * per_u:(rank + 1) * per_u])
    unit_local = torch.zeros(unit_elems, device=dev, dtype=torch.bfloat16)   # stale local copy, filled by the kernel
    w_view = unit_local[:Nw * Kw].view(Nw, Kw)
    b_view = unit_local[Nw * Kw:Nw * Kw + Nw]
    torch.manual_seed(100 + rank)
    x2 = torch.randn(Mx, Kw, device=dev, dtype=torch.bfloat16)
    out = torch.empty(Mx, Nw, device=dev, dtype=torch.bfloat16)
    arrived = torch.zeros(world, device=dev, dtype=torch.int32)
    nch = (per_u * 2 + 32767) // 32768
    torch.cuda.synchronize()
    dist.barrier()
    for gen in (1, 2):  # twice: the arrival counters are monotonic across launches
        if gen == 2:
            unit_local.zero_()
        arena.gemm_allgather(x2, w_view, out, shard_buf=shard16, shard_start=0, per_elems=per_u, unit_dst=unit_local,
                             arrived=arrived, target=gen * nch, bias=b_view, act="gelu_tanh")
        arena.end_step()
        torch.cuda.synchronize()
        ref = torch.nn.functional.gelu(x2.float() @ full_ref[:Nw * Kw].view(Nw, Kw).float().t() + full_ref[Nw * Kw:Nw * Kw + Nw].float(),
                                       approximate="tanh")
        check(f"gemm_allgather[{gen}]", _rel(out, ref) < 2e-2, _rel(out, ref))
        check(f"gemm_allgather_unit[{gen}]", torch.equal(unit_local, full_ref))
        dist.barrier()

    # ---- bandwidth (device-timed, max over ranks)
    if bench:
        from hyperion_b200.parallel.runtime import max_over_ranks

        res = {}
        for dt in (torch.bfloat16, torch.float32):
            buf = bufs[dt]
            for n in (1 << 20, 1 << 23, 1 << 25, N):
                n = min(n, N)
                nbytes = n * buf.element_size()

                def timed(fn, iters=10):
                    for _ in range(3):
                        fn()
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(iters):
                        fn()
                    e1.record()
                    torch.cuda.synchronize()
                    return max_over_ranks(e0.elapsed_time(e1) / iters, dev)

                t_h = timed(lambda: arena.allreduce_avg_(buf, 0, n))
                t_n = timed(lambda: dist.all_reduce(buf[:n]))
                key = f"{str(dt).split('.')[-1]}_{nbytes >> 20}MiB"
                bus = 2 * (world - 1) / world * nbytes / 1e6  # GB/s per ms
                res[key] = {"hyperion_ms": t_h, "nccl_ms": t_n, "hyperion_busbw_GBs": bus / t_h, "nccl_busbw_GBs": bus / t_n}
        info["allreduce_bench"] = res
    info["fails"] = fails
    if out_json:
        os.makedirs(os.path.dirname(out_json) or ".", exist_ok=True)
        gathered = [None] * world
        dist.all_gather_object(gathered, info)
        if rank == 0:
            with open(out_json, "w") as f:
                json.dump(gathered, f, indent=1)
    print(f"[rank {rank}] comm self-test: {'PASS' if not fails else 'FAIL ' + '; '.join(fails)} | {json.dumps({k: v for k, v in info.items() if k != 'fails'})}", flush=True)
    nfail = torch.tensor([len(fails)], device=dev)
    dist.all_reduce(nfail)
    arena.close()
    cleanup()
    return int(nfail.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bench", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    sys.exit(1 if run(a.bench, a.out) else 0)


if __name__ == "__main__":
    main()
