"""CPU tests of the real-corpus preparation path (reference: 02_development/dataset_preparation.ipynb c0, c3, c5): run offline on
in-memory rows, on the archive layout under data/raw, and through the CLI; the oracle for the image transform is torchvision's."""
import json
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent

TEXT = {"train": [" = Title = \n", "\n", "   \n", "the cat sat on the mat \n", "", "a " * 300 + "\n"],
        "validation": ["\n", "the dog \n"], "test": ["the mat \n", " \t \n"]}


def tiny_tokenizer():
    from tokenizers import Tokenizer
    from tokenizers.models import WordLevel
    from tokenizers.pre_tokenizers import Whitespace
    from transformers import PreTrainedTokenizerFast

    vocab = {w: i for i, w in enumerate(["<|endoftext|>", "the", "cat", "sat", "on", "mat", "dog", "a", "=", "Title"])}
    t = Tokenizer(WordLevel(vocab, unk_token="<|endoftext|>"))
    t.pre_tokenizer = Whitespace()
    return PreTrainedTokenizerFast(tokenizer_object=t, eos_token="<|endoftext|>")


def test_wikitext_filter_tokenize_pad_and_save(tmp_path):
    from datasets import load_from_disk

    from hyperion_b200.data import WikiText2TorchDataset, prepare

    tok = tiny_tokenizer()
    assert tok.pad_token is None
    dd = prepare.prepare_wikitext2(tmp_path, source=TEXT, tokenizer=tok, max_length=16, verbose=False)
    assert {s: len(dd[s]) for s in dd} == {"train": 3, "validation": 1, "test": 1}       # blank / whitespace-only lines dropped
    back = load_from_disk(str(tmp_path / "data/processed/wikitext2_tokenized"))
    assert set(back["train"].column_names) == {"input_ids", "attention_mask"}
    ds = WikiText2TorchDataset(back, "train")
    assert ds.ids.shape == (3, 16) and ds.ids.dtype == torch.long
    pad = tok.eos_token_id
    assert tok.pad_token == tok.eos_token                                               # pad falls back to eos, like the notebook
    assert ds.ids[1].tolist() == [1, 2, 3, 4, 1, 5] + [pad] * 10 and ds.mask[1].tolist() == [1] * 6 + [0] * 10
    assert ds.mask[2].tolist() == [1] * 16 and ds.ids[2].tolist() == [7] * 16           # 300 tokens truncated to max_length
    # the same rows through the per-example oracle
    for i, line in enumerate([t for t in TEXT["train"] if t.strip()]):
        enc = tok(line, truncation=True, padding="max_length", max_length=16)
        assert enc["input_ids"] == ds.ids[i].tolist() and enc["attention_mask"] == ds.mask[i].tolist()


def test_wikitext_from_archive_layout_and_hf_rows(tmp_path):
    from datasets import Dataset, DatasetDict

    from hyperion_b200.data import prepare

    raw = tmp_path / "data/raw/wikitext-2-raw"
    raw.mkdir(parents=True)
    for split, f in prepare.WIKITEXT_FILES.items():
        (raw / f).write_text("".join(t if t.endswith("\n") else t + "\n" for t in TEXT[split]), encoding="utf-8")
    a = prepare.prepare_wikitext2(tmp_path, tokenizer=tiny_tokenizer(), max_length=8, download=False, verbose=False)
    b = prepare.prepare_wikitext2(tmp_path / "b", source=DatasetDict({s: Dataset.from_dict({"text": v}) for s, v in TEXT.items()}),
                                  tokenizer=tiny_tokenizer(), max_length=8, verbose=False)
    for s in TEXT:
        assert a[s]["input_ids"] == b[s]["input_ids"] and a[s]["attention_mask"] == b[s]["attention_mask"]


def test_no_network_gives_a_clear_error(tmp_path, monkeypatch):
    from hyperion_b200.data import prepare

    monkeypatch.setenv("HF_HUB_OFFLINE", "1")
    monkeypatch.setenv("HF_DATASETS_OFFLINE", "1")
    with pytest.raises(prepare.DataUnavailable, match="data synth"):
        prepare.prepare_wikitext2(tmp_path, tokenizer=tiny_tokenizer(), download=False)
    with pytest.raises(prepare.DataUnavailable, match="data synth"):
        prepare.prepare_cifar10(tmp_path, download=False)
    with pytest.raises(prepare.DataUnavailable, match="tokenizer"):
        prepare.load_tokenizer(str(tmp_path / "no_such_tokenizer"))


def test_cifar_normalize_filter_matches_torchvision(tmp_path):
    import torchvision.transforms as T

    from hyperion_b200.data import CIFAR10TorchDataset, prepare

    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, (12, 32, 32, 3), dtype=np.uint8)
    imgs[3] = 0                                  # all zero before normalisation -> -1 everywhere after: kept, as in the notebook
    imgs[5] = 128                                # 128/255 is not exactly 0.5: kept as well
    labels = list(range(10)) + [0, 1]
    out = prepare.prepare_cifar10(tmp_path, source={"train": (imgs, labels), "test": (imgs[:2], labels[:2])}, verbose=False)
    assert len(out["train"]) == 12 and len(out["test"]) == 2
    rows = torch.load(tmp_path / "data/processed/cifar10_train.pt", weights_only=False)
    assert isinstance(rows, list) and isinstance(rows[0], tuple) and isinstance(rows[0][1], int)
    tf = T.Compose([T.ToTensor(), T.Normalize((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))])
    for i in (0, 3, 7, 11):
        img, lbl = CIFAR10TorchDataset(rows)[i]
        assert lbl == labels[i] and img.shape == (3, 32, 32) and img.dtype == torch.float32
        torch.testing.assert_close(img, tf(imgs[i]), rtol=0, atol=1e-6)
    # the filter: wrong shapes are dropped (an image that is all zero AFTER normalisation would need every byte = 127.5)
    ragged = [imgs[0], np.zeros((16, 16, 3), np.uint8), imgs[1], np.zeros((32, 32), np.uint8)]
    kept = prepare.normalize_and_filter(ragged, [4, 5, 6, 7])
    assert [l for _, l in kept] == [4, 6]
    assert prepare.normalize_and_filter(np.zeros((0, 32, 32, 3), np.uint8), []) == []
    ppm = (tmp_path / "data/processed/cifar10_sample.ppm").read_bytes()
    assert ppm.startswith(b"P6\n32 32\n255\n") and len(ppm) == 13 + 32 * 32 * 3
    assert np.array_equal(np.frombuffer(ppm[13:], np.uint8).reshape(32, 32, 3), imgs[0])       # un-normalised back to the bytes


def test_verify_and_cli(tmp_path):
    from hyperion_b200.data import prepare, synth

    synth.write_wikitext2_like(tmp_path / "data/processed/wikitext2_tokenized", rows={"train": 30, "validation": 4, "test": 5})
    synth.write_cifar10_like(tmp_path / "data/processed", rows={"train": 6, "test": 3})
    rep = prepare.verify(tmp_path, verbose=False)
    assert rep["wikitext2_rows"] == {"train": 30, "validation": 4, "test": 5} and rep["wikitext2_seq_len"] == 128
    assert len(rep["wikitext2_sample"]["input_ids"]) == 20 and sum(rep["wikitext2_sample"]["attention_mask"]) > 0
    assert 0.0 < rep["wikitext2_pad_fraction"] < 1.0 and rep["matches_reference_rows"] is False
    assert rep["cifar10_train"]["rows"] == 6 and rep["cifar10_train"]["shape"] == (3, 32, 32) and rep["cifar10_test"]["rows"] == 3
    r = subprocess.run([sys.executable, "-m", "hyperion_b200", "data", "verify", "--base_dir", str(tmp_path)], cwd=ROOT,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "wikitext2_rows" in r.stdout and "cifar10_test" in r.stdout
    r = subprocess.run([sys.executable, "-m", "hyperion_b200", "data", "prepare", "--base_dir", str(tmp_path / "x"), "--no-download",
                        "--only", "cifar10"], cwd=ROOT, capture_output=True, text=True, timeout=300,
                       env={**__import__("os").environ, "HF_HUB_OFFLINE": "1"})
    assert r.returncode == 1 and "data synth" in r.stderr
    assert (tmp_path / "x/data/raw").is_dir() and (tmp_path / "x/data/processed").is_dir()      # the directory tree of c0
    json.dumps(rep)                                                                              # report is plain data
