"""Hardware exploration -- the walk-through of the reference's Phase 1/01_hardware_exploration.ipynb.

    python examples/hardware_exploration.py [--out results/benchmarks/hardware] [--quick]

Environment banner, basic-op timings, the matmul TFLOPS sweep per precision (library GEMM and the tcgen05 GEMM of this framework side by
side), the HBM bandwidth sweep, two CSVs and the two figures.  Same as `python -m hyperion_b200 probe`.
"""
from _common import parser

from hyperion_b200.bench import probe


def main(argv=None) -> int:
    ap = parser(__doc__)
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    out = a.out or f"{a.base_dir}/results/benchmarks/hardware"
    args = ["--out", out]
    if a.quick:
        args += ["--sizes", "256,512", "--bw-sizes", "10"]
    return probe.main(args)


if __name__ == "__main__":
    raise SystemExit(main())
