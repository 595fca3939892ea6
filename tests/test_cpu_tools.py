"""CPU tests of the benchmark tools and CLIs: output schemas must equal the reference's."""
import json
import subprocess
import sys
from pathlib import Path

import pandas as pd
import pytest
import torch
import torch.nn as nn

ROOT = Path(__file__).resolve().parent.parent


def test_probe_csv_schemas(tmp_path):
    from hyperion_b200.bench import probe

    assert probe.main(["--out", str(tmp_path), "--sizes", "256", "--bw-sizes", "10"]) == 0
    prec = pd.read_csv(tmp_path / "precision_results.csv")
    assert list(prec.columns)[:4] == ["Size", "Precision", "Time (s)", "TFLOPS"]
    assert set(prec["Precision"]) == {"FP32", "TF32", "FP16", "BF16"}
    bw = pd.read_csv(tmp_path / "bandwidth_results.csv")
    assert list(bw.columns)[:2] == ["Size (M elements)", "Bandwidth (GB/s)"]
    # the notebook's two figures (PNG with matplotlib, SVG without)
    assert any((tmp_path / f"precision_performance.{e}").exists() for e in ("png", "svg"))
    assert any((tmp_path / f"memory_bandwidth.{e}").exists() for e in ("png", "svg"))


def test_benchmark_utils_return_keys(tmp_path):
    from hyperion_b200.bench.utils import (benchmark_forward_pass, benchmark_training_step, compare_precision_formats,
                                           save_benchmark_results)

    keys = {"total_time", "avg_time", "throughput", "memory_usage"}
    net = lambda: nn.Sequential(nn.Linear(16, 32), nn.ReLU(), nn.Linear(32, 16))
    assert set(benchmark_forward_pass(net(), (16,), 4, 3, 1)) == keys
    assert set(benchmark_training_step(net(), (16,), (16,), 4, 3, 1)) == keys
    df = compare_precision_formats(net, (16,), (16,), batch_sizes=(2, 4), precisions=("fp32",), iterations=2, warmup=1)
    assert len(df) == 2 and {"batch_size", "precision", "throughput"} <= set(df.columns)
    # the reference's own call forms (Phase 1/benchmarking.py:12-17, 57-66, 153-158, 223-227)
    m = net()
    x, y = torch.rand(4, 16), torch.rand(4, 16)
    assert set(benchmark_forward_pass(m, x, num_iterations=3, warmup=1)) == keys
    opt = torch.optim.Adam(m.parameters(), lr=3e-3)
    assert set(benchmark_training_step(m, x, y, opt, torch.nn.MSELoss(), num_iterations=3, warmup=1)) == keys
    df2 = compare_precision_formats(net, (16,), batch_sizes=[2], num_iterations=2, warmup=1, precisions=("fp32",))
    assert len(df2) == 1
    p = save_benchmark_results(df2, "ref_form.csv", results_dir=str(tmp_path))
    assert p.endswith("ref_form.csv") and (tmp_path / "ref_form.csv").exists()


def test_baseline_benchmark_columns(tmp_path):
    from hyperion_b200.bench.baseline import COLUMNS, benchmark_model, test_batch_size_scaling as batch_scaling
    from hyperion_b200.models import create_custom_transformer

    r = benchmark_model(create_custom_transformer, (2, 16, 512), (2, 16, 512), 2, 1)
    assert set(COLUMNS) <= set(r)
    assert COLUMNS[:7] == ["Model", "Forward Time (ms)", "Backward Time (ms)", "Optimizer Time (ms)", "Total Time (ms)",
                       "Memory Usage (MB)", "Throughput (samples/s)"]
    tiny = lambda: nn.Sequential(nn.Flatten(), nn.Linear(3 * 8 * 8, 10))
    tiny.__name__ = "tiny"
    df = batch_scaling(tiny, (3, 8, 8), 10, batch_sizes=(1, 2), out_dir=str(tmp_path), iters=2, warmup=1)
    assert list(df.columns) == COLUMNS + ["Batch Size"]
    assert (tmp_path / "scaling" / "tiny_batch_scaling.csv").exists()
    import xml.dom.minidom as minidom

    from hyperion_b200.bench.baseline import visualize_baseline_results, visualize_batch_scaling

    one = pd.DataFrame([r], columns=COLUMNS)
    f1 = visualize_baseline_results(one, str(tmp_path))
    f2 = visualize_batch_scaling(df, "Tiny Net", str(tmp_path))
    assert f1.endswith("visualizations/baseline_benchmarks.svg") and f2.endswith("visualizations/tiny_net_batch_scaling.svg")
    for f in (f1, f2):                                               # 2x2 page of nested charts, panel files removed
        assert len(minidom.parse(f).documentElement.getElementsByTagName("svg")) == 4
    assert sorted(x.name for x in (tmp_path / "visualizations").iterdir()) == ["baseline_benchmarks.svg", "tiny_net_batch_scaling.svg"]


def test_fusion_bench_outputs(tmp_path):
    from hyperion_b200.bench import fusion

    df = fusion.run(str(tmp_path), "fp32", 1, variants=("eager", "hyperion"))
    assert list(df.columns) == ["model", "variant", "time_ms", "mem_gb"]
    assert set(df["model"]) == {"simple_transformer_lm", "resnet18_cifar10"}
    out = tmp_path / "results/benchmarks/compilation"
    assert (out / "compilation_ckpt_benchmark.csv").exists() and (out / "compilation_ckpt_analysis.txt").exists()
    rows = json.loads((out / "compilation_ckpt_benchmark.json").read_text())
    assert set(rows[0]) == {"model", "variant", "time_ms", "mem_gb"}
    assert (out / "compilation_ckpt_speed_mem.png").exists() or (out / "compilation_ckpt_speed.svg").exists()


def test_svg_charts_are_well_formed(tmp_path):
    import xml.dom.minidom as minidom

    from hyperion_b200.utils.svgplot import _ticks, bar_chart, line_chart

    assert _ticks(0, 1800) == [0, 500, 1000, 1500] and _ticks(0, 0.9)[:3] == [0, 0.2, 0.4]
    p = line_chart(tmp_path / "a" / "l.svg", {"ideal": ([1, 2, 4, 8], [1, 2, 4, 8]), "m<1>": ([1, 2, 4, 8], [1, 1.9, None, float("nan")]),
                                               "empty": ([1], [None])}, title="t & t", xlabel="GPUs", ylabel="x")
    d = minidom.parse(str(p))
    assert len(d.getElementsByTagName("polyline")) == 2          # the all-None series is dropped
    assert len(d.getElementsByTagName("circle")) == 4 + 2        # None / NaN points are skipped
    assert "m&lt;1&gt;" in p.read_text() and "stroke-dasharray" in p.read_text()
    b = bar_chart(tmp_path / "b.svg", ["x", "y"], {"eager": [1.0, 2.0], "hyperion": [0.5, None]}, title="ms")
    d = minidom.parse(str(b))
    bars = [r for r in d.getElementsByTagName("rect") if r.getAttribute("fill").startswith("#") and r.getAttribute("width") != "12"]
    assert len(bars) == 3
    assert max(float(r.getAttribute("height")) for r in bars) <= 310
    line_chart(tmp_path / "e.svg", {})                               # degenerate inputs still give a valid file
    minidom.parse(str(tmp_path / "e.svg"))


def test_bench_reference_arm_reports_unavailable():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d


def test_cli_help_and_data_synth(tmp_path):
    r = subprocess.run([sys.executable, "-m", "hyperion_b200"], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 2 and "probe" in r.stdout
    r = subprocess.run([sys.executable, "-m", "hyperion_b200", "data", "synth", "--small", "--base_dir", str(tmp_path)],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-500:]
    assert (tmp_path / "data/processed/cifar10_train.pt").exists()
    assert (tmp_path / "data/processed/wikitext2_tokenized/dataset_dict.json").exists()


def test_header_declares_every_exported_entry_point():
    """include/hyperion_b200.h is the contract: every `HYP_API` symbol in csrc/ must be declared there."""
    import re

    hdr = (ROOT / "include/hyperion_b200.h").read_text()
    missing = []
    for cu in list((ROOT / "csrc").glob("*.cu")) + list((ROOT / "csrc/common").glob("*.cu")):
        for m in re.finditer(r"HYP_API\s+[\w\s\*]+?\b(hyp_\w+)\s*\(", cu.read_text()):
            if not re.search(r"\b" + m.group(1) + r"\s*\(", hdr):
                missing.append(m.group(1))
    assert not missing, missing


def test_reference_arm_environment_and_unavailable_line(tmp_path, capsys, monkeypatch):
    """bench.py --impl reference: configs the reference cannot run print one `unavailable` JSON line and exit 0; for the
    ones it can, the environment the unmodified trainers need (tokenizer dir, on-disk dataset, import stand-ins) is built
    and the installed reference module imports from baseline/_ref."""
    import json
    import sys
    import types

    from baseline import reference_arm as ra

    rc = ra.run(types.SimpleNamespace(config="gpt2_fsdp", steps=2, warmup=1))
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert rc == 0 and line["impl"] == "reference" and "unavailable" in line and line["same_config"] is False
    assert set(line["reference_configs"]) == {"lm_ddp", "lm_fsdp", "cifar_ddp"}
    if not (ra.REF_DIR / "distributed_utils.py").exists():
        pytest.skip("baseline/_ref not installed on this machine (baseline/install_reference.sh)")
    monkeypatch.setenv("RANK", "0")
    ra._prepare(tmp_path, "lm_ddp", 8, 128)
    assert (tmp_path / "gpt2" / "tokenizer.json").exists() and (tmp_path / ".ready").exists()
    assert (tmp_path / "data/processed/wikitext2_tokenized/dataset_dict.json").exists()
    made = ra._shim_modules()
    assert "peft" in sys.modules and "matplotlib.pyplot" in sys.modules
    monkeypatch.syspath_prepend(str(ra.REF_DIR))
    monkeypatch.setenv("HF_HUB_OFFLINE", "1")
    import importlib

    ref = importlib.import_module("distributed_utils")
    assert {"train_language_model_ddp", "train_language_model_fsdp", "train_cifar_model_ddp"} <= set(dir(ref))
    monkeypatch.chdir(tmp_path)
    from transformers import AutoTokenizer

    tok = AutoTokenizer.from_pretrained("gpt2", use_fast=True)
    tok.pad_token = tok.pad_token or tok.eos_token
    assert tok.vocab_size == 50257 and tok.pad_token_id == 50256
    for name in made:
        sys.modules.pop(name, None)


def test_distributed_utils_drop_in_module():
    """`import distributed_utils` of the reference keeps working: same twelve public names, the reference's argument order."""
    import importlib.util
    import inspect

    spec = importlib.util.spec_from_file_location("hyperion_distributed_utils_dropin", ROOT / "distributed_utils.py")
    du = importlib.util.module_from_spec(spec)            # by path: the reference arm imports the reference's module of that name
    spec.loader.exec_module(du)
    names = ["WikiText2TorchDataset", "CIFAR10TorchDataset", "SimpleTransformerLM", "_local_gpu", "setup", "cleanup",
             "train_language_model_ddp", "train_cifar_model_ddp", "train_language_model_fsdp", "train_llama_fsdp",
             "create_scaling_report", "run_scaling_experiment"]
    assert sorted(du.__all__) == sorted(names) and all(callable(getattr(du, n)) for n in names)
    for fn in ("train_language_model_ddp", "train_cifar_model_ddp", "train_language_model_fsdp"):
        assert list(inspect.signature(getattr(du, fn)).parameters)[:4] == ["rank", "world_size", "epochs", "base_dir"]
    p = inspect.signature(du.train_llama_fsdp).parameters
    assert {"epochs", "base_dir", "hf_token", "model_id", "lora", "batch_size", "progress_every"} <= set(p)
    assert list(inspect.signature(du.setup).parameters)[:3] == ["rank", "world_size", "backend"]
    assert list(inspect.signature(du.run_scaling_experiment).parameters)[:5] == ["model_type", "gpu_counts", "epochs", "base_dir", "hf_token"]
    cmds = du.run_scaling_experiment("llama", [1, 2], 1, "/tmp/x", "tok", dry_run=True)
    assert len(cmds) == 2 and "--hf_token=tok" in cmds[1] and "--model=llama" in cmds[1] and "--nproc-per-node=2" in cmds[1]
    m = du.SimpleTransformerLM(100, 32, 2, 1)             # positional like distributed_utils.py:76
    assert m(torch.randint(0, 100, (2, 8))).shape == (2, 8, 100)
    assert du._local_gpu(3) == 0 or torch.cuda.is_available()


def test_benchmarking_drop_in_module():
    import importlib.util

    spec = importlib.util.spec_from_file_location("hyperion_benchmarking_dropin", ROOT / "benchmarking.py")
    bm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bm)
    r = bm.benchmark_forward_pass(nn.Linear(8, 8), torch.rand(4, 8), num_iterations=2, warmup=1)
    assert set(r) == {"total_time", "avg_time", "throughput", "memory_usage"}
    assert sorted(bm.__all__) == ["benchmark_forward_pass", "benchmark_training_step", "compare_precision_formats", "save_benchmark_results"]


def test_every_environment_switch_is_documented():
    """TROUBLESHOOTING.md (the reference README's section 9: environment knobs, common issues) lists every HYPERION_* variable the code reads."""
    import re

    doc = (ROOT / "TROUBLESHOOTING.md").read_text()
    used = set()
    files = [*ROOT.glob("*.py"), *(ROOT / "hyperion_b200").rglob("*.py"), *(ROOT / "csrc").rglob("*.cu*"), *(ROOT / "scripts").glob("*.sh"),
             *(ROOT / "scripts").glob("*.py"), *(ROOT / "baseline").glob("*.py")]
    for f in files:
        used |= set(re.findall(r"HYPERION_[A-Z0-9_]+", f.read_text()))
    used -= {"HYPERION_B200_H"}                      # include guard
    missing = sorted(v for v in used if v not in doc)
    assert not missing, f"undocumented environment switches: {missing}"
    for section in ("## Environment switches", "## Common issues", "## Reproducibility"):
        assert section in doc


def test_baseline_md_matrix_equals_the_committed_json_lines():
    """Every row of the round-2 matrix in BASELINE.md is what scripts/matrix_table.py prints from profiles/bench/r2/*.json, and the
    headline ratios quoted under it follow from the same files."""
    r = subprocess.run([sys.executable, str(ROOT / "scripts/matrix_table.py"), str(ROOT / "profiles/bench/r2")], capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    doc = (ROOT / "BASELINE.md").read_text()
    rows = r.stdout.strip().splitlines()
    assert len(rows) > 20 and all(line in doc for line in rows)

    def val(n, cfg, arm):
        return json.loads((ROOT / f"profiles/bench/r2/matrix_n{n}_{cfg}_{arm}.json").read_text().splitlines()[0])["value"]

    for cfg, base, n, quoted in (("lm_ddp", "reference", 1, "2.95x"), ("lm_ddp", "reference", 8, "3.06x"), ("lm_fsdp", "reference", 8, "4.46x"),
                                 ("cifar_ddp", "reference", 1, "3.54x"), ("gpt2_fsdp", "torch", 1, "2.78x"), ("gpt2_fsdp", "torch", 8, "3.15x"),
                                 ("resnet50_ddp", "torch", 1, "1.26x"), ("vit_ddp", "torch", 1, "1.84x"), ("llama_lora", "torch", 8, "4.22x")):
        ratio = val(n, cfg, "hyperion") / val(n, cfg, base)
        assert abs(ratio - float(quoted[:-1])) < 0.011 and quoted in doc, (cfg, n, ratio, quoted)
    for f in (ROOT / "profiles/bench/r2").glob("matrix_n*_*.json"):           # every line carries a clocks record or says unavailable
        j = json.loads(f.read_text().splitlines()[0])
        assert "unavailable" in j or "clocks" in j, f.name


def test_ctypes_signatures_agree_with_the_c_header():
    """Every entry point's ctypes signature (ops/_lib.py table + the declare() calls of flash.py / conv.py / symm.py) has the argument count
    and the argument kinds (pointer, int, long long, float) of its prototype in include/hyperion_b200.h -- a drifted signature would
    corrupt arguments silently at call time."""
    import ast
    import re

    from hyperion_b200.ops import _lib

    header = (ROOT / "include/hyperion_b200.h").read_text()
    header = re.sub(r"/\*.*?\*/", " ", header, flags=re.S)
    header = re.sub(r"//[^\n]*", " ", header)
    protos = {}
    for ret, name, args in re.findall(r"\b([A-Za-z_][\w\s\*]*?)\b(hyp_\w+)\s*\(([^;{]*?)\)\s*;", header):
        codes = ""
        for a in [x.strip() for x in args.split(",") if x.strip() and x.strip() != "void"]:
            if "*" in a or "cudaStream_t" in a:
                codes += "p"
            elif "long long" in a or "size_t" in a or "int64_t" in a:
                codes += "l"
            elif "float" in a:
                codes += "f"
            elif re.search(r"\b(int|unsigned)\b", a):
                codes += "i"
            else:
                raise AssertionError(f"{name}: cannot classify parameter {a!r}")
        protos[name] = codes
    assert len(protos) > 40

    def classify(args: str) -> str:
        codes = ""
        for a in [x.strip() for x in args.split(",") if x.strip() and x.strip() != "void"]:
            codes += ("p" if "*" in a or "cudaStream_t" in a else "l" if "long long" in a or "size_t" in a or "int64_t" in a
                      else "f" if "float" in a else "i")
        return codes

    # the .cu files do not include the public header, so the compiler does not compare definition and prototype: do it here
    defined = {}
    for cu in (ROOT / "csrc").rglob("*.cu"):
        src = re.sub(r"/\*.*?\*/", " ", cu.read_text(), flags=re.S)
        src = re.sub(r"//[^\n]*", " ", src)
        for name, args in re.findall(r"HYP_API\s+[\w\s\*]+?\b(hyp_\w+)\s*\(([^{;]*?)\)\s*\{", src):
            defined[name] = classify(args)
    assert len(defined) > 40
    for name, codes in sorted(defined.items()):
        assert name in protos and codes == protos[name], f"{name}: definition {codes} != header {protos.get(name)}"

    declared = dict(_lib._SIGS)
    for rel in ("hyperion_b200/ops/flash.py", "hyperion_b200/ops/conv.py", "hyperion_b200/parallel/symm.py", "hyperion_b200/ops/native.py",
                "hyperion_b200/ops/flash_bwd.py", "hyperion_b200/ops/functional.py", "hyperion_b200/optim.py", "hyperion_b200/amp.py"):
        tree = ast.parse((ROOT / rel).read_text())
        for node in ast.walk(tree):
            if isinstance(node, ast.Call) and getattr(node.func, "attr", getattr(node.func, "id", "")) == "declare" and len(node.args) >= 2 \
                    and isinstance(node.args[0], ast.Constant):
                declared[node.args[0].value] = eval(compile(ast.Expression(node.args[1]), rel, "eval"), {})
    assert len(declared) > 35
    for name, sig in sorted(declared.items()):
        assert name in protos, f"{name} is declared for ctypes but has no prototype in the header"
        assert sig == protos[name], f"{name}: ctypes {sig} ({len(sig)} args) != header {protos[name]} ({len(protos[name])} args)"
