"""NVTX ranges around the phases of a training step (forward / backward / comm / optimizer), so an Nsight Systems or
``ncu --nvtx --nvtx-include`` capture can be cut by phase.  Enabled with ``HYPERION_NVTX=1``
(off by default: a range is two host calls per phase per step, which the eager path of a 1.9 ms step can see).
The reference has no tracing hooks at all; this file has no counterpart there."""
from __future__ import annotations

import contextlib
import os

import torch

ENABLED = os.environ.get("HYPERION_NVTX", "0") == "1"


@contextlib.contextmanager
def range(name: str):  # noqa: A001 - mirrors torch.cuda.nvtx.range
    if ENABLED and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def mark(name: str) -> None:
    if ENABLED and torch.cuda.is_available():
        torch.cuda.nvtx.mark(name)
