"""NVTX ranges around stages, fits and aggregations (``B2_NVTX=1``), so an ``nsys`` / ``ncu --nvtx`` capture of a
federation shows which node is in which stage next to the kernels.  Off by default: a range is two driver calls per
stage and the control plane runs thousands of stages per minute.  Not in the reference (its only profiling hook is
yappi in the example script, examples/mnist.py:264-297)."""

from __future__ import annotations

import contextlib
import os

ENABLED = os.environ.get("B2_NVTX", "0") == "1"


@contextlib.contextmanager
def nvtx_range(name: str):
    if not ENABLED:
        yield
        return
    try:
        import torch

        ok = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        ok = False
    if not ok:
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()
