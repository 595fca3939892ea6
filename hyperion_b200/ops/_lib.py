"""ctypes loader for ``libhyperion_b200.so`` (the C ABI in ``include/hyperion_b200.h``).

The library is built in-tree by :mod:`hyperion_b200.build`.  On a machine with a
CUDA device the ops in this package *require* it: a missing or unloadable library
raises instead of silently falling back to eager PyTorch (the GPU tests must
exercise the sm_100a kernels).  On a CPU-only machine (`torch.cuda.is_available()
== False`) :func:`available` is False and callers use their PyTorch reference path.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libhyperion_b200.so"
_lib = None
_load_error: str | None = None

_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
_CODES = {"p": _P, "i": _I, "l": _L, "f": _F}

# name -> argument codes (see the header for the meaning of every argument)
_SIGS = {
    "hyp_gemm": "pppiiillliiiipiplipiiffip",
    "hyp_gemm_dualk": "pppiiillliiii" + "ppill" + "pi" + "pli" + "fp",
    "hyp_stream_add": "ppplip",
    "hyp_stream_copy": "pplp",
    "hyp_colsum": "ppiilip",
    "hyp_colsum_cast": "ppiilippp",
    "hyp_swiglu_fwd": "ppplip",
    "hyp_swiglu_bwd": "ppppplip",
    "hyp_cast": "ppliifp",
    "hyp_cast_dscale": "ppliifpp",
    "hyp_norm_fwd": "ppppppppiifiip",
    "hyp_norm_fwd_drop": "ppppppppiifii" + "pif" + "p",
    "hyp_dropout": "pplii" + "pif" + "p",
    "hyp_norm_bwd": "pppppppppliiiiip",
    "hyp_norm_bwd_fused": "ppppppppppipliiiip",
    "hyp_adamw": "ppppplii" + "fffff" + "i" + "f" + "ppppp",
    "hyp_sumsq": "plipp",
    "hyp_scaler_note": "ppp",
    "hyp_scaler_update": "ppppfffp",
    "hyp_ce_fwd_bwd": "plpiilpppfiip",
    "hyp_ce_finalize": "pppp",
    "hyp_embedding_fwd": "ppplilip",
    "hyp_embedding_bwd": "ppplillip",
    "hyp_rope": "piiiilllpfiip",
    "hyp_conv2d_wgrad": "ppp" + "iiiiiiiii" + "ip",
    "hyp_conv2d_fwd": "ppp" + "iiiiiiiii" + "ppp" + "ii" + "p",
}

DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def lib_path() -> Path:
    return _LIB_PATH


def _load():
    global _lib, _load_error
    if _lib is not None or _load_error is not None:
        return _lib
    if not _LIB_PATH.exists():
        _load_error = f"{_LIB_PATH} not found; run `python -m hyperion_b200.build`"
        return None
    try:
        lib = ctypes.CDLL(str(_LIB_PATH))
    except OSError as e:  # pragma: no cover
        _load_error = f"cannot load {_LIB_PATH}: {e}"
        return None
    lib.hyp_last_error.restype = ctypes.c_char_p
    lib.hyp_norm_bwd_scratch_bytes.restype = _L
    lib.hyp_norm_bwd_scratch_bytes.argtypes = [_I, _I]
    lib.hyp_norm_bwd_fused_ws_bytes.restype = _L
    lib.hyp_norm_bwd_fused_ws_bytes.argtypes = [_I]
    for name, sig in _SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.restype = _I
        fn.argtypes = [_CODES[c] for c in sig]
    # optional entry points declared by later modules
    _lib = lib
    return lib


def raw():
    """The ctypes.CDLL handle (None when unavailable)."""
    return _load()


def available() -> bool:
    """True when the native kernels can run: library loaded AND a CUDA device present."""
    if os.environ.get("HYPERION_DISABLE_NATIVE") == "1":
        return False
    return torch.cuda.is_available() and _load() is not None


def require() -> ctypes.CDLL:
    lib = _load()
    if lib is None:
        raise RuntimeError(f"hyperion_b200 native library unavailable: {_load_error}")
    return lib


def declare(name: str, sig: str, restype=_I) -> None:
    """Register the signature of an entry point added by another module."""
    lib = require()
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = [_CODES[c] for c in sig]
    _SIGS[name] = sig


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


_launches = 0
# kernels launched per entry-point call (everything not listed launches exactly one)
_KERNELS_PER_CALL = {"hyp_norm_bwd": 2, "hyp_ipc_alloc": 0, "hyp_ipc_open": 0, "hyp_ipc_close": 0, "hyp_ipc_free": 0}


def launch_count() -> int:
    """Number of sm_100a kernels of this library launched so far (bench.py reports the delta)."""
    return _launches


def add_launches(n: int) -> None:
    """Account for kernels launched by a CUDA-graph replay (they do not pass through :func:`call`)."""
    global _launches
    _launches += n


def call(name: str, *args) -> None:
    """Call an int-returning entry point and raise with hyp_last_error() on failure."""
    global _launches
    lib = require()
    rc = getattr(lib, name)(*args)
    _launches += _KERNELS_PER_CALL.get(name, 1)
    if rc != 0:
        msg = lib.hyp_last_error().decode(errors="replace")
        raise RuntimeError(f"{name} failed (rc={rc}): {msg}")
