"""Multi-GPU tests (need >= 2 GPUs on the box): comm-kernel self-test vs NCCL and fused-vs-NCCL training parity.
Each spawns torchrun with one rank per GPU; kernels that wait on peers are never run as separate ranks on one GPU."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = Path(__file__).resolve().parent.parent


def _torchrun(n, module, *args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1",
           f"--nproc-per-node={n}", "-m", module, *args]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:   # keep the whole log where a gpurun call brings it back
        out = ROOT / "gpurun_out"
        out.mkdir(exist_ok=True)
        (out / ("multi_fail_" + "_".join(a.strip("-") for a in args) + ".log")).write_text(r.stdout + "\n---- stderr ----\n" + r.stderr)
    rank0 = "\n".join(l for l in r.stderr.splitlines() if "[rank1]" not in l)
    assert r.returncode == 0, (r.stdout[-2000:], rank0[-4000:])
    return r.stdout


def _n():
    return min(torch.cuda.device_count(), 8)


def test_comm_kernels_match_nccl():
    out = _torchrun(_n(), "hyperion_b200.parallel.selftest")
    assert "PASS" in out and "FAIL" not in out


@pytest.mark.parametrize("strategy,precision", [("ddp", "bf16"), ("ddp", "fp16"), ("fsdp", "bf16")])
def test_fused_training_matches_nccl_path(strategy, precision):
    out = _torchrun(_n(), "hyperion_b200.parallel.parity", "--strategy", strategy, "--precision", precision)
    assert "PASS" in out and "FAIL" not in out


def test_full_shard_mode_matches_nccl_path():
    """reshard_after_forward=True (FULL_SHARD: slots, gather before forward and again before backward) on the fused NVLink
    kernels vs the params-resident NCCL path."""
    out = _torchrun(_n(), "hyperion_b200.parallel.parity", "--strategy", "fsdp", "--precision", "bf16", "--reshard")
    assert "PASS" in out and "FAIL" not in out


@pytest.mark.parametrize("reshard", [False, True])
def test_block_units_gather_inside_their_first_gemm(reshard):
    """GPT-2-shaped model with one FSDP unit per Block: fused path = all-gather warps inside each block's c_attn GEMM plus
    stand-alone gathers for the root unit (and, with --reshard, FULL_SHARD slots) vs the NCCL path."""
    args = ["--strategy", "fsdp", "--precision", "bf16", "--model", "gpt2"] + (["--reshard"] if reshard else [])
    import os

    os.environ["HYPERION_FSDP_AG_BLOCKS"] = "1"   # inherited by the ranks
    try:
        out = _torchrun(_n(), "hyperion_b200.parallel.parity", *args)
    finally:
        os.environ.pop("HYPERION_FSDP_AG_BLOCKS", None)
    assert "PASS" in out and "FAIL" not in out


@pytest.mark.parametrize("extra", [[], ["--reshard"], ["--model", "gpt2", "--reshard"]])
def test_no_kernel_reads_a_parameter_before_its_gather_landed(extra):
    """Race-detection build (HYPERION_COMM_POISON=1): gather destinations are NaN-filled right before each gather kernel,
    so any consumer that is not ordered after its gather event -- resident units, FULL_SHARD slots being reused, block
    units -- produces a NaN loss and fails the parity run."""
    import os

    os.environ["HYPERION_COMM_POISON"] = "1"
    try:
        out = _torchrun(_n(), "hyperion_b200.parallel.parity", "--strategy", "fsdp", "--precision", "bf16", *extra)
    finally:
        os.environ.pop("HYPERION_COMM_POISON", None)
    assert "PASS" in out and "FAIL" not in out
