"""Mixed precision -- the walk-through of the reference's 02_development/mixed_precision.ipynb.

    python examples/mixed_precision.py --base_dir /tmp/h [--epochs 25] [--precision fp16|bf16] [--quick]

The LM and ResNet-18 trained under mixed precision (fp16: 16-bit module + fp32 master weights + dynamic loss scale whose inf check runs on
the device; bf16: no scaler), checkpoints saved as {"model_state_dict", "scaler_state_dict"} to
data/processed/{simple_transformer_lm,resnet18_cifar10}_amp.pt, then the notebook's profile: 10 LM batches timed with the peak memory, here
for fp32, fp16 and bf16 next to each other.
"""
from pathlib import Path

import torch
from _common import device_banner, parser

from hyperion_b200.data import synth
from hyperion_b200.models import NotebookTransformerLM, resnet18
from hyperion_b200.train.single import profile_amp_training, train_cifar_model, train_language_model


def main(argv=None) -> int:
    ap = parser(__doc__)
    ap.add_argument("--epochs", type=int, default=25)
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--profile-batches", type=int, default=10)
    a = ap.parse_args(argv)
    dev = device_banner()
    proc = Path(a.base_dir) / "data" / "processed"
    kw = dict(base_dir=a.base_dir, synthetic=not a.real_data, device=dev, precision=a.precision)
    vocab, pad, bs = (50257, 50256, 32) if not a.quick else (512, 511, 8)
    if a.quick:
        a.epochs, a.profile_batches = min(a.epochs, 2), 2
        kw.update(rows=64, steps_per_epoch=2)

    print(f"Training language model on WikiText-2 with AMP/{a.precision} ({a.epochs} epochs)...")
    train_language_model(NotebookTransformerLM(vocab), a.epochs, batch_size=bs, vocab_size=vocab, pad_id=pad,
                         save_path=proc / "simple_transformer_lm_amp.pt", **kw)
    print(f"Training ResNet-18 on CIFAR-10 with AMP/{a.precision} ({a.epochs} epochs)...")
    train_cifar_model(resnet18(num_classes=10), a.epochs, batch_size=bs, save_path=proc / "resnet18_cifar10_amp.pt", **kw)
    ck = torch.load(proc / "simple_transformer_lm_amp.pt")
    if isinstance(ck, dict) and "model_state_dict" in ck:
        print("AMP-trained models and scaler state saved:", sorted(ck), "scaler:", ck["scaler_state_dict"])
    else:
        print("no loss scaler in use (bf16, or a CPU-only run): bare state_dicts saved.")

    print(f"\nProfiling training (language model, {a.profile_batches} batches)...")
    rows = []
    for prec in ("fp32", "fp16", "bf16"):
        batches = [synth.token_batch(bs, 128, vocab, "cpu", seed=i, pad_id=pad, pad_fraction=0.3)[0] for i in range(a.profile_batches)]
        r = profile_amp_training(NotebookTransformerLM(vocab), batches, precision=prec, pad_id=pad, device=dev)
        rows.append((prec, r["time_s"], r["peak_mb"]))
    print("\nprecision  time (s)  peak (MB)")
    for prec, t, m in rows:
        print(f"{prec:9s}  {t:8.3f}  {m:9.2f}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
