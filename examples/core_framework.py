"""Core framework -- the walk-through of the reference's 02_development/core_framework.ipynb.

    python examples/core_framework.py --base_dir /tmp/h [--epochs 25] [--real-data] [--quick]

Environment check; load the tokenised WikiText-2 and the CIFAR-10 tensors; the 2-layer Transformer LM (AdamW 2e-4, cross entropy with the
pad id ignored, next-token shift) and ResNet-18 (AdamW 1e-3, accuracy) trained in fp32 on one GPU; bare state_dicts saved to
data/processed/{simple_transformer_lm,resnet18_cifar10}.pt; both files loaded back into fresh models (what the later notebooks do).
"""
from pathlib import Path

import numpy as np
import torch
from _common import device_banner, parser

from hyperion_b200.models import NotebookTransformerLM, resnet18
from hyperion_b200.train.single import train_cifar_model, train_language_model


def main(argv=None) -> int:
    ap = parser(__doc__)
    ap.add_argument("--epochs", type=int, default=25)
    a = ap.parse_args(argv)
    print("torch:", torch.__version__, "| numpy:", np.__version__)
    dev = device_banner()
    proc = Path(a.base_dir) / "data" / "processed"
    kw = dict(base_dir=a.base_dir, synthetic=not a.real_data, device=dev)
    if a.quick:
        a.epochs = min(a.epochs, 2)
        kw.update(rows=64, steps_per_epoch=2)
    vocab = 50257 if not a.quick else 512

    print(f"Training language model on WikiText-2 ({a.epochs} epochs, sanity check)...")
    lm = NotebookTransformerLM(vocab)
    lm_losses = train_language_model(lm, a.epochs, batch_size=32 if not a.quick else 8, vocab_size=vocab,
                                     pad_id=50256 if not a.quick else 511, save_path=proc / "simple_transformer_lm.pt", **kw)
    print(f"Training ResNet-18 on CIFAR-10 ({a.epochs} epochs, sanity check)...")
    net = resnet18(num_classes=10)
    hist = train_cifar_model(net, a.epochs, batch_size=32 if not a.quick else 8, save_path=proc / "resnet18_cifar10.pt", **kw)
    print(f"Models saved after {a.epochs} epochs.")

    # the checkpoints are bare state_dicts with the reference's parameter names: a fresh model takes them strictly
    sd = torch.load(proc / "simple_transformer_lm.pt")
    missing = NotebookTransformerLM(vocab).load_state_dict(sd, strict=True)
    sd2 = torch.load(proc / "resnet18_cifar10.pt")
    missing2 = resnet18(num_classes=10).load_state_dict(sd2, strict=True)
    print(f"Reloaded: LM {len(sd)} tensors {missing}, ResNet-18 {len(sd2)} tensors {missing2}")
    print(f"LM loss {lm_losses[0]:.4f} -> {lm_losses[-1]:.4f}; CIFAR loss {hist[0][0]:.4f} -> {hist[-1][0]:.4f}, acc {hist[-1][1]:.2f}%")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
