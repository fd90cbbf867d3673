"""NVTX ranges for timelines (Nsight Systems / CUPTI consumers).

Off by default: with ``B200TRAIN_NVTX=1`` every model unit's forward / backward, every collective of
the peer / NCCL communicators and the optimizer phases open a named range (``torch.cuda.nvtx``), so a
trace shows ``unit3.fwd``, ``unit3.bwd``, ``rs.unit3`` ... on top of the kernel rows. The reference has
no tracing hooks (only ``LocalTimer`` phases and py-spy).
"""
from __future__ import annotations

import contextlib
import os

import torch

ENABLED = os.environ.get("B200TRAIN_NVTX", "0") == "1" and torch.cuda.is_available()


def push(name: str):
    if ENABLED:
        torch.cuda.nvtx.range_push(name)


def pop():
    if ENABLED:
        torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def range(name: str):  # noqa: A001 - mirrors torch.cuda.nvtx.range
    if not ENABLED:
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()
