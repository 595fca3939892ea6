"""Builds ``libhyperion_b200.so`` (every CUDA kernel of the framework) in-tree.

    python -m hyperion_b200.build [--force] [--verbose]

Each ``csrc/*.cu`` is compiled for sm_100a only
(``-gencode arch=compute_100a,code=sm_100a -lineinfo``) into ``build/obj`` and the
objects are linked into ``hyperion_b200/lib/libhyperion_b200.so``.  nvcc
cross-compiles without a GPU, so this also is the "does it build" check that
``__graft_entry__.build()`` runs.  The reference has no native code and no build
system; this file has no counterpart there.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OBJ = ROOT / "build" / "obj"
LIB_DIR = ROOT / "hyperion_b200" / "lib"
LIB = LIB_DIR / "libhyperion_b200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "--expt-relaxed-constexpr", "-Xptxas", "-v", "-I", str(CSRC), "-I", str(ROOT / "include"),
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set $NVCC")


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu")) + sorted((CSRC / "common").glob("*.cu"))


def _headers_mtime() -> float:
    hdrs = list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return max((h.stat().st_mtime for h in hdrs), default=0.0)


def _compile(src: Path, verbose: bool) -> tuple[Path, str]:
    obj = OBJ / (src.stem + ".o")
    cmd = [_nvcc(), *ARCH_FLAGS, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    (OBJ / (src.stem + ".ptxas.txt")).write_text(r.stderr)
    return obj, r.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    srcs = _sources()
    hdr_m = _headers_mtime()
    todo, objs = [], []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        objs.append(o)
        if force or not o.exists() or o.stat().st_mtime < max(s.stat().st_mtime, hdr_m):
            todo.append(s)
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            list(ex.map(lambda s: _compile(s, verbose), todo))
    if todo or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        # shared cudart: the process must have ONE runtime (torch's), so that the current
        # device and streams mean the same thing on both sides of the C ABI.
        cmd = [_nvcc(), *ARCH_FLAGS, "-shared", "-cudart", "shared", "-o", str(LIB), *map(str, objs),
               "-Xlinker", "-rpath,/usr/local/cuda/lib64"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))


if __name__ == "__main__":
    main()
