"""``python -m hyperion_b200 <command>``: probe | baseline | fusion-bench | data | comm-check | scaling | build."""
from __future__ import annotations

import sys


def main() -> int:
    cmds = ["probe", "baseline", "fusion-bench", "data", "comm-check", "scaling", "build"]
    if len(sys.argv) < 2 or sys.argv[1] not in cmds:
        print("usage: python -m hyperion_b200 {" + ",".join(cmds) + "} [options]")
        return 2
    cmd, argv = sys.argv[1], sys.argv[2:]
    if cmd == "probe":
        from .bench.probe import main as m
        return m(argv)
    if cmd == "baseline":
        from .bench.baseline import main as m
        return m(argv)
    if cmd == "fusion-bench":
        from .bench.fusion import main as m
        return m(argv)
    if cmd == "build":
        from .build import build
        print(build(force="--force" in argv))
        return 0
    if cmd == "data":
        import argparse

        from .data import synth

        ap = argparse.ArgumentParser(prog="hyperion data")
        ap.add_argument("what", choices=["synth", "prepare", "verify"],
                        help="synth: seeded stand-ins in the real formats; prepare: the real corpora (data/raw or a download); "
                             "verify: reload and print what is under data/processed")
        ap.add_argument("--base_dir", default=".")
        ap.add_argument("--seed", type=int, default=0)
        ap.add_argument("--small", action="store_true", help="synth: 1/50 of the row counts (quick check)")
        ap.add_argument("--only", choices=["wikitext2", "cifar10"], default=None, help="prepare: one corpus instead of both")
        ap.add_argument("--tokenizer", default="gpt2", help="prepare: tokenizer name or local directory")
        ap.add_argument("--no-download", action="store_true", help="prepare: use data/raw only")
        a = ap.parse_args(argv)
        if a.what == "verify":
            from .data import prepare

            return 0 if prepare.verify(a.base_dir) else 1
        if a.what == "prepare":
            from .data import prepare

            try:
                if a.only in (None, "wikitext2"):
                    prepare.prepare_wikitext2(a.base_dir, tokenizer=a.tokenizer, download=not a.no_download)
                if a.only in (None, "cifar10"):
                    prepare.prepare_cifar10(a.base_dir, download=not a.no_download)
            except prepare.DataUnavailable as e:
                print(f"error: {e}", file=sys.stderr)
                return 1
            prepare.verify(a.base_dir)
            return 0
        rows_w = {k: max(8, v // 50) for k, v in synth.WIKITEXT2_ROWS.items()} if a.small else None
        rows_c = {k: max(8, v // 50) for k, v in synth.CIFAR10_ROWS.items()} if a.small else None
        synth.write_wikitext2_like(f"{a.base_dir}/data/processed/wikitext2_tokenized", seed=a.seed, rows=rows_w)
        synth.write_cifar10_like(f"{a.base_dir}/data/processed", seed=a.seed, rows=rows_c)
        print(f"wrote synthetic WikiText-2 / CIFAR-10 shaped data under {a.base_dir}/data/processed")
        return 0
    if cmd == "comm-check":
        import os

        if "RANK" not in os.environ:
            print("run under torchrun: python -m torch.distributed.run --standalone --nproc-per-node N -m hyperion_b200 comm-check")
            return 2
        from .parallel.selftest import run
        return 1 if run("--bench" in argv, None) else 0
    if cmd == "scaling":
        import argparse

        from .utils.scaling import run_scaling_experiment

        ap = argparse.ArgumentParser(prog="hyperion scaling")
        ap.add_argument("--model", default="language_ddp")
        ap.add_argument("--gpus", default="1,2,4,8")
        ap.add_argument("--epochs", type=int, default=3)
        ap.add_argument("--base_dir", default=".")
        ap.add_argument("--steps", type=int, default=None)
        a = ap.parse_args(argv)
        extra = [f"--steps={a.steps}"] if a.steps else []
        run_scaling_experiment(a.model, [int(g) for g in a.gpus.split(",")], a.epochs, a.base_dir, extra)
        return 0
    return 2


if __name__ == "__main__":
    sys.exit(main())
