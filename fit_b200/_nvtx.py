"""NVTX ranges around the phases of the hot path (the reference leans on the Lightning profiler / nsys for this).

Off by default -- a range costs two Python calls per phase -- and switched on with FITB200_NVTX=1 (read once at import) for `nsys` / `ncu
--nvtx` sessions: `ncu --nvtx --nvtx-include "fit/block/"` then profiles the kernels of the transformer blocks only.  Ranges pushed during a
CUDA-graph capture mark the capture, not the replays."""
from __future__ import annotations

import contextlib
import os

import torch

ENABLED = os.environ.get("FITB200_NVTX", "0") == "1"


@contextlib.contextmanager
def _range(name: str):
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


_NULL = contextlib.nullcontext()


def rng(name: str):
    """``with rng("fit/block/3"):`` -- an NVTX range when FITB200_NVTX=1, a no-op context otherwise."""
    return _range(name) if ENABLED else _NULL
