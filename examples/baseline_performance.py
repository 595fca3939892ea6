"""Baseline model benchmarks -- the walk-through of the reference's Phase 1/baseline_performance.ipynb.

    python examples/baseline_performance.py [--precision fp32|tf32|fp16|bf16] [--quick]

Forward / backward / optimizer time, peak memory and throughput of ResNet-50, ViT-B/16, the fallback CNN the reference published as "ViT"
and the custom Transformer (Adam 1e-3, MSE against random targets), the ResNet-50 batch-size sweep 1..64, the CSVs and the two 2x2 figure
pages.  Same as `python -m hyperion_b200 baseline`.
"""
from _common import parser

from hyperion_b200.bench import baseline


def main(argv=None) -> int:
    ap = parser(__doc__)
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32", "fp16", "bf16"])
    a = ap.parse_args(argv)
    args = ["--out", f"{a.base_dir}/results/benchmarks", "--precision", a.precision]
    if a.quick:
        args += ["--batch", "2", "--iters", "2", "--warmup", "1", "--no-scaling"]
    return baseline.main(args)


if __name__ == "__main__":
    raise SystemExit(main())
