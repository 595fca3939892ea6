"""make_notebooks.py [out_dir]: writes the Jupyter notebooks under ``notebooks/`` (nbformat 4 JSON, no jupyter needed to generate).

One notebook per notebook of the reference, same order of steps, built on this framework's API:

    01_hardware_exploration   Phase 1/01_hardware_exploration.ipynb
    baseline_performance      Phase 1/baseline_performance.ipynb
    dataset_preparation       02_development/dataset_preparation.ipynb
    core_framework            02_development/core_framework.ipynb
    mixed_precision           02_development/mixed_precision.ipynb
    memory_optimization       02_development/memory_optimization.ipynb

The cell sources live here as plain Python so they are linted and diffed like code; ``tests/test_cpu_examples.py`` checks that the committed
``.ipynb`` files equal what this script writes and executes every code cell with ``HYPERION_NOTEBOOK_QUICK=1`` on the CPU.
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

SETUP = '''import os, sys
REPO = os.environ.get("HYPERION_REPO", os.path.abspath(".." if os.path.basename(os.getcwd()) == "notebooks" else "."))
sys.path.insert(0, REPO)
BASE_DIR = os.environ.get("HYPERION_BASE_DIR", REPO)            # data/processed and results/ live under it
QUICK = os.environ.get("HYPERION_NOTEBOOK_QUICK", "0") == "1"   # a few steps on a few rows: seconds, also on a CPU
import torch
device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
print("Using device:", device, torch.cuda.get_device_name(0) if device.type == "cuda" else "")'''

NOTEBOOKS: dict[str, list[tuple[str, str]]] = {}

NOTEBOOKS["01_hardware_exploration"] = [
    ("md", "# Hardware exploration (B200)\nEnvironment, basic-op timings, matmul TFLOPS per precision (library GEMM next to this framework's "
           "tcgen05 GEMM), HBM bandwidth, CSVs and figures.  Command-line form: `python -m hyperion_b200 probe`."),
    ("code", SETUP),
    ("code", '''from hyperion_b200.bench import probe
from hyperion_b200.memory import get_gpu_memory
env = probe.environment()
print(f"Python {env['python']} | torch {env['torch']} | CUDA {env['cuda']}")
for d in env["devices"]:
    print(f"  GPU {d['index']}: {d['name']} ({d['sms']} SMs, {d['memory_gb']:.1f} GB, cc {d['cc']})")
print("Memory allocated / reserved (MB):", get_gpu_memory())'''),
    ("code", '''print(probe.test_gpu_operations(device))          # tensor creation, 5000^2 fp32 matmul, sin + cos (median ms)'''),
    ("md", "## Precision formats\n`TFLOPS = 2 n^3 / t`; FP32 is IEEE fp32 (TF32 off), TF32 / FP16 / BF16 run on the tensor cores."),
    ("code", '''sizes = (256, 512) if QUICK or device.type == "cpu" else (1024, 2048, 4096, 8192)
prec = probe.test_precision_formats(device, sizes)
prec'''),
    ("md", "## Memory bandwidth\n`z = x + y` in fp32, 12 bytes per element."),
    ("code", '''bw = probe.test_memory_bandwidth(device, (10,) if QUICK or device.type == "cpu" else (10, 20, 50, 100, 200, 500))
bw'''),
    ("code", '''out = os.path.join(BASE_DIR, "results/benchmarks/hardware")
os.makedirs(out, exist_ok=True)
prec.to_csv(os.path.join(out, "precision_results.csv"), index=False)
bw.to_csv(os.path.join(out, "bandwidth_results.csv"), index=False)
probe.plot_results(prec, bw, out)'''),
]

NOTEBOOKS["baseline_performance"] = [
    ("md", "# Baseline model benchmarks\nForward / backward / optimizer time, peak memory and throughput of ResNet-50, ViT-B/16, the fallback CNN "
           "the reference published as 'ViT' and the custom Transformer; ResNet-50 batch-size sweep.  Command-line form: "
           "`python -m hyperion_b200 baseline`."),
    ("code", SETUP),
    ("code", '''from hyperion_b200.bench import baseline
kw = dict(batch=2, iters=2, warmup=1) if QUICK else dict(batch=32, iters=50, warmup=10)
results = baseline.run_baseline_benchmarks(precision="fp32", out_dir=os.path.join(BASE_DIR, "results/benchmarks"), **kw)
results'''),
    ("code", '''baseline.visualize_baseline_results(results, os.path.join(BASE_DIR, "results"))'''),
    ("md", "## Batch-size scaling (ResNet-50)"),
    ("code", '''from hyperion_b200.models import create_resnet50
sizes = (1, 2) if QUICK else (1, 2, 4, 8, 16, 32, 64)
it = dict(iters=2, warmup=1) if QUICK else dict(iters=20, warmup=5)
scaling = baseline.test_batch_size_scaling(create_resnet50, batch_sizes=sizes, out_dir=os.path.join(BASE_DIR, "results/benchmarks"), **it)
scaling'''),
    ("code", '''baseline.visualize_batch_scaling(scaling, "ResNet-50", os.path.join(BASE_DIR, "results"))'''),
]

NOTEBOOKS["dataset_preparation"] = [
    ("md", "# Dataset preparation\nWikiText-2 (blank lines dropped, GPT-2 tokenizer, pad = eos, length 128) and CIFAR-10 (normalised to [-1, 1], "
           "list of `(image, label)`), written in the formats every trainer reads.  Without the corpora (no network, nothing under `data/raw`) "
           "the same formats are written from a seeded generator.  Command-line form: `python -m hyperion_b200 data prepare | synth | verify`."),
    ("code", SETUP),
    ("code", '''from hyperion_b200.data import prepare, synth
dirs = prepare.make_dirs(BASE_DIR)
print("Directory structure created:")
for p in dirs.values():
    print(" ", p)'''),
    ("md", "## WikiText-2"),
    ("code", '''try:
    if QUICK:
        raise prepare.DataUnavailable("quick run")
    prepare.prepare_wikitext2(BASE_DIR, tokenizer="gpt2")
except prepare.DataUnavailable as e:
    print("real corpus not available:", str(e)[:120], "...\\nwriting the synthetic stand-in")
    rows = {k: max(8, v // 200) for k, v in synth.WIKITEXT2_ROWS.items()} if QUICK else None
    synth.write_wikitext2_like(dirs["processed"] / "wikitext2_tokenized", rows=rows)'''),
    ("md", "## CIFAR-10"),
    ("code", '''try:
    if QUICK:
        raise prepare.DataUnavailable("quick run")
    prepare.prepare_cifar10(BASE_DIR)
except prepare.DataUnavailable as e:
    print("real corpus not available:", str(e)[:120], "...\\nwriting the synthetic stand-in")
    rows = {k: max(8, v // 500) for k, v in synth.CIFAR10_ROWS.items()} if QUICK else None
    synth.write_cifar10_like(dirs["processed"], rows=rows)'''),
    ("md", "## Verify"),
    ("code", '''report = prepare.verify(BASE_DIR)'''),
]

NOTEBOOKS["core_framework"] = [
    ("md", "# Core framework\nLoad the prepared data, build the 2-layer Transformer LM and ResNet-18, train both on one GPU in fp32 "
           "(AdamW 2e-4 / 1e-3, 25 epochs in the reference), save bare `state_dict`s."),
    ("code", SETUP),
    ("code", '''import numpy as np
print("torch:", torch.__version__, "| numpy:", np.__version__)
if torch.cuda.is_available():
    print("CUDA device count:", torch.cuda.device_count(), "| current:", torch.cuda.current_device(), "|", torch.cuda.get_device_name(0))'''),
    ("md", "## Data\n`data/processed` if `dataset_preparation` was run, synthetic rows of the same shape otherwise."),
    ("code", '''from hyperion_b200.data import DeviceResidentLoader
from hyperion_b200.train.common import load_cifar_tensors, load_wikitext_tensors
have = os.path.exists(os.path.join(BASE_DIR, "data/processed/wikitext2_tokenized"))
rows = 64 if QUICK else None
ids, mask = load_wikitext_tensors(BASE_DIR, device, synthetic=not have, rows=rows)
images, labels = load_cifar_tensors(BASE_DIR, device, synthetic=not have, rows=rows)
print("WikiText-2:", tuple(ids.shape), "sample input_ids:", ids[0, :20].tolist())
print("CIFAR-10:", tuple(images.shape), "label:", int(labels[0]))
loader = DeviceResidentLoader((ids,), 32, shuffle=True)          # whole split on the device, no host work per step
print("DataLoaders ready:", len(loader), "LM batches per epoch")'''),
    ("md", "## Models"),
    ("code", '''from hyperion_b200.models import NotebookTransformerLM, resnet18
vocab_size = int(ids.max()) + 1 if QUICK else 50257
model_lm = NotebookTransformerLM(vocab_size, emb_dim=256, n_heads=4, n_layers=2, max_len=128)
model_cifar = resnet18(num_classes=10)
print(sum(p.numel() for p in model_lm.parameters()), "LM parameters;", sum(p.numel() for p in model_cifar.parameters()), "ResNet-18 parameters")'''),
    ("md", "## Train"),
    ("code", '''from hyperion_b200.train.single import train_cifar_model, train_language_model
proc = os.path.join(BASE_DIR, "data/processed")
epochs = 2 if QUICK else 25
kw = dict(base_dir=BASE_DIR, synthetic=not have, device=device, **(dict(rows=64, steps_per_epoch=2) if QUICK else {}))
lm_losses = train_language_model(model_lm, epochs, vocab_size=vocab_size, batch_size=8 if QUICK else 32,
                                 save_path=os.path.join(proc, "simple_transformer_lm.pt"), **kw)
history = train_cifar_model(model_cifar, epochs, batch_size=8 if QUICK else 32, save_path=os.path.join(proc, "resnet18_cifar10.pt"), **kw)
print("Models saved after", epochs, "epochs.")'''),
    ("md", "## Reload\nThe files are bare `state_dict`s with the reference's parameter names."),
    ("code", '''print(NotebookTransformerLM(vocab_size).load_state_dict(torch.load(os.path.join(proc, "simple_transformer_lm.pt"))))
print(resnet18(num_classes=10).load_state_dict(torch.load(os.path.join(proc, "resnet18_cifar10.pt"))))'''),
]

NOTEBOOKS["mixed_precision"] = [
    ("md", "# Mixed precision\nThe same two trainings under fp16 (16-bit module, fp32 master weights, dynamic loss scale with the inf check on "
           "the device) or bf16, checkpoints with the scaler state, and the 10-batch profile for fp32 / fp16 / bf16."),
    ("code", SETUP),
    ("code", '''from hyperion_b200.data import synth
from hyperion_b200.models import NotebookTransformerLM, resnet18
from hyperion_b200.train.single import profile_amp_training, train_cifar_model, train_language_model
proc = os.path.join(BASE_DIR, "data/processed")
precision = "fp16"
vocab, pad, bs = (512, 511, 8) if QUICK else (50257, 50256, 32)
epochs = 2 if QUICK else 25
kw = dict(base_dir=BASE_DIR, device=device, precision=precision, **(dict(rows=64, steps_per_epoch=2) if QUICK else {}))'''),
    ("code", '''print(f"Training language model on WikiText-2 with AMP ({epochs} epochs)...")
train_language_model(NotebookTransformerLM(vocab), epochs, batch_size=bs, vocab_size=vocab, pad_id=pad,
                     save_path=os.path.join(proc, "simple_transformer_lm_amp.pt"), **kw)
print(f"Training ResNet-18 on CIFAR-10 with AMP ({epochs} epochs)...")
train_cifar_model(resnet18(num_classes=10), epochs, batch_size=bs, save_path=os.path.join(proc, "resnet18_cifar10_amp.pt"), **kw)'''),
    ("code", '''ck = torch.load(os.path.join(proc, "simple_transformer_lm_amp.pt"))
print(sorted(ck) if "model_state_dict" in ck else "no loss scaler in use (bf16, or a CPU-only run): bare state_dict saved")'''),
    ("md", "## Profile: time and peak memory of 10 LM batches"),
    ("code", '''n = 2 if QUICK else 10
for prec in ("fp32", "fp16", "bf16"):
    batches = [synth.token_batch(bs, 128, vocab, "cpu", seed=i, pad_id=pad, pad_fraction=0.3)[0] for i in range(n)]
    print(prec, profile_amp_training(NotebookTransformerLM(vocab), batches, precision=prec, pad_id=pad, device=device))'''),
]

NOTEBOOKS["memory_optimization"] = [
    ("md", "# Memory optimisation\nAllocated / reserved / peak memory around a forward pass and around one training batch, with and without "
           "activation checkpointing (`use_checkpoint=True` on the LM, `checkpoint_resnet_blocks` on the ResNet)."),
    ("code", SETUP),
    ("code", '''from hyperion_b200.data import synth
from hyperion_b200.memory import MemoryStats, checkpoint_resnet_blocks, print_memory
from hyperion_b200.models import NotebookTransformerLM, resnet18
from hyperion_b200.train.single import train_one_batch
proc = os.path.join(BASE_DIR, "data/processed")
vocab, pad, bs = (512, 511, 4) if QUICK else (50257, 50256, 32)
model_lm, model_cifar = NotebookTransformerLM(vocab), resnet18(num_classes=10)
for m, f in ((model_lm, "simple_transformer_lm.pt"), (model_cifar, "resnet18_cifar10.pt")):
    if os.path.exists(os.path.join(proc, f)):
        sd = torch.load(os.path.join(proc, f))
        if next(iter(sd.values())).shape == next(iter(m.state_dict().values())).shape:
            m.load_state_dict(sd); print("loaded", f)
model_lm, model_cifar = model_lm.to(device), model_cifar.to(device)
input_ids = synth.token_batch(bs, 128, vocab, device, seed=0, pad_id=pad, pad_fraction=0.3)[0]
images = torch.stack([im for im, _ in synth.cifar10_like_list(bs, seed=0)]).to(device)
mem = MemoryStats(device)'''),
    ("md", "## Baseline forward"),
    ("code", '''def forward_report(title, model, x):
    print(title); mem.reset_peak(); print_memory("Before forward pass:")
    with torch.no_grad():
        model.eval()(x)
    print_memory("After forward pass:")
forward_report("Baseline memory usage: Language model", model_lm, input_ids[:, :-1])
forward_report("Baseline memory usage: CIFAR-10 model", model_cifar, images)'''),
    ("md", "## Activation checkpointing"),
    ("code", '''model_lm_ckpt = NotebookTransformerLM(vocab, use_checkpoint=True).to(device)
model_lm_ckpt.load_state_dict(model_lm.state_dict())
model_cifar_ckpt = checkpoint_resnet_blocks(model_cifar)
forward_report("Language model with activation checkpointing", model_lm_ckpt, input_ids[:, :-1])
forward_report("ResNet-18 with activation checkpointing", model_cifar_ckpt, images)'''),
    ("md", "## One training batch\nThe peak is where checkpointing shows (a forward under `no_grad` keeps no activations either way)."),
    ("code", '''for name, model in (("no checkpointing", model_lm), ("with checkpointing", model_lm_ckpt)):
    if device.type == "cuda":
        torch.cuda.empty_cache()
    r = train_one_batch(model, input_ids, pad_id=pad, device=device)
    print(f"{name}: before {r['allocated_before_mb']:.2f} MB | after {r['allocated_after_mb']:.2f} MB | peak {r['peak_mb']:.2f} MB | loss {r['loss']:.4f}")'''),
]


def notebook(cells: list[tuple[str, str]]) -> dict:
    out = []
    for kind, src in cells:
        lines = src.split("\n")
        source = [l + "\n" for l in lines[:-1]] + [lines[-1]]
        if kind == "md":
            out.append({"cell_type": "markdown", "metadata": {}, "source": source})
        else:
            out.append({"cell_type": "code", "execution_count": None, "metadata": {}, "outputs": [], "source": source})
    return {"cells": out, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                       "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}


def write_all(out_dir) -> list[Path]:
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    paths = []
    for name, cells in NOTEBOOKS.items():
        p = out_dir / f"{name}.ipynb"
        p.write_text(json.dumps(notebook(cells), indent=1) + "\n")
        paths.append(p)
    return paths


if __name__ == "__main__":
    for p in write_all(sys.argv[1] if len(sys.argv) > 1 else ROOT / "notebooks"):
        print(p)
