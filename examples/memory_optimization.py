"""Memory optimisation -- the walk-through of the reference's 02_development/memory_optimization.ipynb.

    python examples/memory_optimization.py --base_dir /tmp/h [--quick]

Loads the checkpoints core_framework.py saved (fresh weights if they are absent), prints allocated / reserved / PEAK memory around a forward
of the LM and ResNet-18, builds the activation-checkpointed variants (the LM's `use_checkpoint=True` with the weights copied over,
`checkpoint_resnet_blocks` on the ResNet), repeats the forward, then trains one batch with and without checkpointing.  The reference's
forward comparison runs under no_grad and so shows no difference by construction; the one-batch comparison reports the peak, which is where
checkpointing shows.
"""
from pathlib import Path

import torch
from _common import device_banner, parser

from hyperion_b200.data import synth
from hyperion_b200.memory import MemoryStats, checkpoint_resnet_blocks, print_memory
from hyperion_b200.models import NotebookTransformerLM, resnet18
from hyperion_b200.train.single import train_one_batch


def _load(model, path):
    if Path(path).exists():
        ck = torch.load(path)
        model.load_state_dict(ck.get("model_state_dict", ck) if isinstance(ck, dict) else ck)
        print(f"loaded {path}")
    return model


def main(argv=None) -> int:
    a = parser(__doc__).parse_args(argv)
    dev = device_banner()
    proc = Path(a.base_dir) / "data" / "processed"
    vocab, pad, bs = (50257, 50256, 32) if not a.quick else (512, 511, 4)
    lm = _load(NotebookTransformerLM(vocab), proc / "simple_transformer_lm.pt").to(dev)
    net = _load(resnet18(num_classes=10), proc / "resnet18_cifar10.pt").to(dev)
    ids = synth.token_batch(bs, 128, vocab, dev, seed=0, pad_id=pad, pad_fraction=0.3)[0]
    images = synth.cifar10_like_list(bs, seed=0)
    images = torch.stack([im for im, _ in images]).to(dev)
    mem = MemoryStats(dev)

    def forward_report(title, model, x):
        print(title)
        mem.reset_peak()
        print_memory("Before forward pass:")
        with torch.no_grad():
            model.eval()(x)
        print_memory("After forward pass:")

    forward_report("Baseline memory usage: Language model", lm, ids[:, :-1])
    forward_report("\nBaseline memory usage: CIFAR-10 model", net, images)

    lm_ckpt = NotebookTransformerLM(vocab, use_checkpoint=True).to(dev)
    lm_ckpt.load_state_dict(lm.state_dict())
    print("\nCheckpointed language model ready.")
    net_ckpt = checkpoint_resnet_blocks(net)
    print("Checkpointed ResNet-18 ready.")
    forward_report("\nMemory usage: Language model with activation checkpointing", lm_ckpt, ids[:, :-1])
    forward_report("\nMemory usage: ResNet-18 with activation checkpointing", net_ckpt, images)

    out = {}
    for name, model in (("no checkpointing", lm), ("with checkpointing", lm_ckpt)):
        if dev.type == "cuda":
            torch.cuda.empty_cache()
        print(f"\nTraining one batch ({name}):")
        r = train_one_batch(model, ids, pad_id=pad, device=dev)
        print(f"Before: {r['allocated_before_mb']:.2f} MB | After: {r['allocated_after_mb']:.2f} MB | Peak: {r['peak_mb']:.2f} MB")
        print(f"Loss: {r['loss']:.4f}")
        out[name] = r
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
