"""Shared by the walk-throughs in this directory: repo root on ``sys.path``, the common flags, the device banner."""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def parser(doc: str) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=doc, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--base_dir", default=".", help="project directory (data/processed and results/ live under it)")
    ap.add_argument("--quick", action="store_true", help="a few steps on a few rows: finishes in seconds, also on a CPU")
    ap.add_argument("--real-data", action="store_true", help="read data/processed/* written from the real datasets instead of synthetic rows")
    return ap


def device_banner():
    import torch

    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    print("Using device:", dev, f"({torch.cuda.get_device_name(0)})" if dev.type == "cuda" else "")
    return dev
