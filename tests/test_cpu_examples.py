"""The walk-throughs under examples/ (one per notebook of the reference) run end to end in --quick mode on a CPU and leave the
reference's artefacts behind (dataset_preparation, core_framework, mixed_precision, memory_optimization, the two Phase 1
notebooks)."""
import importlib
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent


def run_example(name, argv):
    sys.path.insert(0, str(ROOT / "examples"))
    try:
        mod = importlib.import_module(name)
        return mod.main(argv)
    finally:
        sys.path.remove(str(ROOT / "examples"))


def test_notebook_walkthroughs_in_order(tmp_path, capsys):
    base = ["--base_dir", str(tmp_path), "--quick"]
    assert run_example("dataset_preparation", base + ["--synthetic"]) == 0
    out = capsys.readouterr().out
    assert "Directory structure created" in out and "wikitext2_rows" in out and "cifar10_train" in out
    assert (tmp_path / "data/raw").is_dir() and (tmp_path / "data/processed/wikitext2_tokenized/dataset_dict.json").exists()

    assert run_example("core_framework", base) == 0
    out = capsys.readouterr().out
    assert out.count("[LM] Epoch") == 2 and out.count("[CIFAR] Epoch") == 2 and out.count("All keys matched") == 2
    sd = torch.load(tmp_path / "data/processed/simple_transformer_lm.pt")
    assert "embedding.weight" in sd and "fc.weight" in sd and all(isinstance(v, torch.Tensor) for v in sd.values())   # bare state_dict
    assert "layer4.1.bn2.running_var" in torch.load(tmp_path / "data/processed/resnet18_cifar10.pt")

    assert run_example("mixed_precision", base + ["--precision", "bf16"]) == 0
    out = capsys.readouterr().out
    assert (tmp_path / "data/processed/simple_transformer_lm_amp.pt").exists() and (tmp_path / "data/processed/resnet18_cifar10_amp.pt").exists()
    assert out.count("AMP Training Time for 2 batches") == 3 and "precision  time (s)  peak (MB)" in out

    assert run_example("memory_optimization", base) == 0
    out = capsys.readouterr().out
    assert out.count("loaded ") == 2                      # the checkpoints of core_framework.py
    assert "Checkpointed language model ready." in out and "Checkpointed ResNet-18 ready." in out
    losses = [float(l.split(":")[1]) for l in out.splitlines() if l.startswith("Loss:")]
    assert len(losses) == 2 and abs(losses[0] - losses[1]) < 1e-4      # same weights, same batch: checkpointing changes memory, not math


def test_phase1_walkthroughs(tmp_path, capsys):
    assert run_example("hardware_exploration", ["--base_dir", str(tmp_path), "--quick"]) == 0
    hw = tmp_path / "results/benchmarks/hardware"
    assert (hw / "precision_results.csv").exists() and (hw / "bandwidth_results.csv").exists()
    assert any(hw.glob("precision_performance.*")) and any(hw.glob("memory_bandwidth.*"))
    assert run_example("baseline_performance", ["--base_dir", str(tmp_path), "--quick"]) == 0
    assert (tmp_path / "results/benchmarks/baseline/model_benchmarks.csv").exists()
    assert (tmp_path / "results/visualizations/baseline_benchmarks.svg").exists()


def test_checkpoint_resnet_blocks_on_a_torchvision_model():
    import torchvision

    from hyperion_b200.memory import checkpoint_resnet_blocks

    torch.manual_seed(0)
    net = torchvision.models.resnet18(num_classes=10)
    x = torch.randn(2, 3, 32, 32)
    ref = net(x)
    ref.sum().backward()
    g_ref = net.conv1.weight.grad.clone()
    net.zero_grad()
    ck = checkpoint_resnet_blocks(net)
    out = ck(x)
    out.sum().backward()
    torch.testing.assert_close(out, ref)
    torch.testing.assert_close(net.conv1.weight.grad, g_ref)
    with torch.no_grad():
        torch.testing.assert_close(ck(x), ref)


def test_notebooks_are_in_sync_and_run(tmp_path, monkeypatch, capsys):
    """notebooks/*.ipynb (one per notebook of the reference): valid nbformat-4 JSON, equal to what scripts/make_notebooks.py writes, and
    every code cell executes in order with HYPERION_NOTEBOOK_QUICK=1 on the CPU."""
    import importlib.util
    import json

    spec = importlib.util.spec_from_file_location("make_notebooks", ROOT / "scripts/make_notebooks.py")
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    fresh = {p.name: p.read_text() for p in mk.write_all(tmp_path / "nb")}
    names = ["01_hardware_exploration", "baseline_performance", "dataset_preparation", "core_framework", "mixed_precision",
             "memory_optimization"]
    assert sorted(fresh) == sorted(n + ".ipynb" for n in names)
    monkeypatch.setenv("HYPERION_NOTEBOOK_QUICK", "1")
    monkeypatch.setenv("HYPERION_BASE_DIR", str(tmp_path))
    monkeypatch.setenv("HYPERION_REPO", str(ROOT))
    for n in ["dataset_preparation", "core_framework", "mixed_precision", "memory_optimization", "01_hardware_exploration",
              "baseline_performance"]:
        committed = (ROOT / "notebooks" / f"{n}.ipynb").read_text()
        assert committed == fresh[n + ".ipynb"], f"notebooks/{n}.ipynb is stale: run python scripts/make_notebooks.py"
        nb = json.loads(committed)
        assert nb["nbformat"] == 4 and nb["cells"][0]["cell_type"] == "markdown"
        ns: dict = {"__name__": f"nb_{n}"}
        for cell in nb["cells"]:
            assert set(cell) >= {"cell_type", "metadata", "source"}
            if cell["cell_type"] == "code":
                assert cell["outputs"] == [] and cell["execution_count"] is None          # committed without outputs
                exec(compile("".join(cell["source"]), f"{n}.ipynb", "exec"), ns)
    out = capsys.readouterr().out
    assert "Directory structure created" in out and "Models saved after 2 epochs." in out and "with checkpointing" in out
    assert (tmp_path / "data/processed/simple_transformer_lm.pt").exists() and (tmp_path / "data/processed/resnet18_cifar10_amp.pt").exists()
    assert (tmp_path / "results/benchmarks/hardware/precision_results.csv").exists()
    assert (tmp_path / "results/visualizations/resnet-50_batch_scaling.svg").exists()
