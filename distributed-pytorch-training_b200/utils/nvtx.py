"""Optional NVTX ranges.  The reference has no tracing at all (train_ddp.py has
only ``time.time()`` around each batch, :196,224); torch DDP wraps its forward in a
``record_function`` range (distributed.py:1694).  Here the phases of a step are named ranges
that show up in any NVTX-aware tool; they are compiled out to a no-op unless enabled with
``--nvtx`` / ``DPT_NVTX=1`` so the hot loop pays nothing by default.
"""
from __future__ import annotations

import contextlib
import os

import torch

_enabled = os.environ.get("DPT_NVTX", "0") == "1"


def enable(on: bool = True) -> None:
    global _enabled
    _enabled = bool(on)


def enabled() -> bool:
    return _enabled and torch.cuda.is_available()


@contextlib.contextmanager
def range(name: str):  # noqa: A001 - mirrors torch.cuda.nvtx.range
    if not enabled():
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def mark(name: str) -> None:
    if enabled():
        torch.cuda.nvtx.mark(name)
