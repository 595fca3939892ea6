"""Dataset preparation -- the walk-through of the reference's 02_development/dataset_preparation.ipynb.

    python examples/dataset_preparation.py --base_dir /data/hyperion            # real corpora (data/raw or a download)
    python examples/dataset_preparation.py --base_dir /tmp/h --synthetic         # seeded stand-ins in the same on-disk formats
    python examples/dataset_preparation.py --base_dir /tmp/h --synthetic --quick

Steps: directory tree; WikiText-2 -> filter blank lines -> GPT-2 tokenizer (pad = eos, length 128) -> save_to_disk; CIFAR-10 ->
normalise -> filter -> torch.save of list[(image, label)]; reload and print the checks.
"""
from _common import parser

from hyperion_b200.data import prepare, synth


def main(argv=None) -> int:
    ap = parser(__doc__)
    ap.add_argument("--synthetic", action="store_true", help="no corpora needed: write seeded data in the same formats")
    ap.add_argument("--tokenizer", default="gpt2")
    ap.add_argument("--no-download", action="store_true")
    a = ap.parse_args(argv)
    d = prepare.make_dirs(a.base_dir)
    print("Directory structure created:")
    for p in d.values():
        print(" ", p)
    if a.synthetic or a.quick:
        rows_w = {k: max(8, v // 200) for k, v in synth.WIKITEXT2_ROWS.items()} if a.quick else None
        rows_c = {k: max(8, v // 500) for k, v in synth.CIFAR10_ROWS.items()} if a.quick else None
        synth.write_wikitext2_like(d["processed"] / "wikitext2_tokenized", rows=rows_w)
        synth.write_cifar10_like(d["processed"], rows=rows_c)
        print("Synthetic WikiText-2 / CIFAR-10 shaped data written to", d["processed"])
    else:
        try:
            prepare.prepare_wikitext2(a.base_dir, tokenizer=a.tokenizer, download=not a.no_download)
            prepare.prepare_cifar10(a.base_dir, download=not a.no_download)
        except prepare.DataUnavailable as e:
            print("error:", e)
            return 1
    print("\nVerification (reloaded from disk):")
    rep = prepare.verify(a.base_dir)
    return 0 if rep else 1


if __name__ == "__main__":
    raise SystemExit(main())
