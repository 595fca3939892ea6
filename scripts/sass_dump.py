"""Full SASS listings of the shipping kernels (one instantiation per kernel family) -> profiles/sass/<kernel>.sass, plus an
index with the mnemonic evidence per kernel (UTCHMMA / UTCQMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load,
SYNCS = mbarrier, LDGMC / multimem, RED/ATOM) and the register / spill line from ptxas.

    python scripts/sass_dump.py            # after python -m hyperion_b200.build
"""
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "hyperion_b200/lib/libhyperion_b200.so"
OUT = ROOT / "profiles/sass"

# (file stem, regex on the demangled name): the instantiation that the BASELINE configs actually launch
WANT = [
    ("gemm2_kernel_plain_bn256", r"gemm::gemm2_kernel<0, 0, 256>"),
    ("gemm2_kernel_bias_gelu_bn256", r"gemm::gemm2_kernel<1, 3, 256>"),
    ("gemm2_kernel_dact_bn256", r"gemm::gemm2_kernel<2, 3, 256>"),
    ("gemm2_kernel_splitk_bn256", r"gemm::gemm2_kernel<4, 0, 256>"),
    ("gemm_kernel_allgather_bn128", r"gemm::gemm_kernel<128, 0, 0, true>"),
    ("flash_fwd_kernel_bf16_d64", r"attn::flash_fwd_kernel<__nv_bfloat16, 64, false>"),
    ("flash_fwd_kernel_bf16_d64_dropout", r"attn::flash_fwd_kernel<__nv_bfloat16, 64, true>"),
    ("flash_fwd_kernel_bf16_d128", r"attn::flash_fwd_kernel<__nv_bfloat16, 128, false>"),
    ("flash_bwd64_kernel_bf16", r"attn::flash_bwd64_kernel<__nv_bfloat16, false>"),
    ("flash_bwd64_kernel_bf16_dropout", r"attn::flash_bwd64_kernel<__nv_bfloat16, true>"),
    ("flash_bwd_kernel_bf16_d128", r"attn::flash_bwd_kernel<__nv_bfloat16, 128, false>"),
    ("conv_fwd_kernel_128", r"conv::conv_fwd_kernel<128>"),
    ("conv_wgrad_kernel_128", r"conv::conv_wgrad_kernel<128>"),
    ("norm_fwd_kernel_bf16_d1024", r"norm::norm_fwd_kernel<__nv_bfloat16, 128, 1, 128, false>"),
    ("norm_fwd_kernel_bf16_d1024_dropout", r"norm::norm_fwd_kernel<__nv_bfloat16, 128, 1, 128, true>"),
    ("norm_bwd_kernel_bf16_d1024", r"norm::norm_bwd_kernel<__nv_bfloat16, 128, 1, 512>"),
    ("ce_fwd_bwd_reg_kernel_bf16", r"hyp::ce_fwd_bwd_reg_kernel<__nv_bfloat16, 13>"),
    ("adamw_kernel_bf16", r"hyp::adamw_kernel<__nv_bfloat16, __nv_bfloat16>"),
    ("dropout_kernel_bf16", r"ew::dropout_kernel<__nv_bfloat16>"),
    ("colsum8_kernel_bf16", r"hyp::colsum8_kernel<__nv_bfloat16>"),
    ("bn_apply_kernel_bf16", r"bn::apply_kernel<__nv_bfloat16>"),
    ("comm_allreduce_kernel", r"comm::allreduce_kernel"),
    ("comm_reduce_scatter_kernel", r"comm::reduce_scatter_kernel"),
    ("comm_allgather_kernel", r"comm::allgather_kernel"),
]
EVIDENCE = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "LDGMC", "MULTIMEM", "REDG",
            "RED", "ATOM", "FFMA2", "FMUL2", "MUFU.EX2", "LDG", "STG", "IMAD.HI", "IMAD.WIDE"]


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", sass)
    funcs = {}
    for p in parts[1:]:
        name, _, body = p.partition("\n")
        funcs[name.strip()] = body
    dem = subprocess.run(["cu++filt"], input="\n".join(funcs), capture_output=True, text=True).stdout.splitlines()
    by_dem = dict(zip(dem, funcs))
    index = ["# kernel file | mangled name | instructions | evidence mnemonics (count)", ""]
    for stem, pat in WANT:
        norm = lambda t: re.sub(r"\((?:int|bool|unsigned int)\)", "", t).replace("true", "1").replace("false", "0").replace(" ", "")
        hit = [d for d in by_dem if norm(pat) in norm(d)]
        if not hit:
            index.append(f"{stem}: NOT FOUND ({pat})")
            continue
        d = sorted(hit, key=len)[0]
        body = funcs[by_dem[d]]
        # keep address + instruction, drop the hex encodings (second column and the continuation lines): 3x smaller
        lean = "\n".join(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln) for ln in body.splitlines()
                         if not re.match(r"^\s*/\* 0x[0-9a-f]+ \*/\s*$", ln))
        (OUT / f"{stem}.sass").write_text(f"// {d}\n// {by_dem[d]}\n{lean}\n")
        ops = re.findall(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", body, flags=re.M)
        counts = {}
        for e in EVIDENCE:
            n = sum(1 for o in ops if o.startswith(e))
            if n:
                counts[e] = n
        index.append(f"{stem}.sass | {by_dem[d]} | {len(ops)} | " + ", ".join(f"{k} {v}" for k, v in counts.items()))
    (OUT / "INDEX.txt").write_text("\n".join(index) + "\n")
    print("\n".join(index))


if __name__ == "__main__":
    main()
