"""Kernel dispatch policy.

CUDA tensors always go to the hand-written sm_100a kernels in ``libb200train.so``; if the library
cannot be loaded on a GPU machine the call raises (no silent PyTorch fallback). CPU tensors (unit
tests, gloo multi-process tests, the chapter-01 CPU plumbing config) run the PyTorch reference math
that the kernels are validated against.
"""
from __future__ import annotations

import os

import torch

from .. import _C

_FORCE_REFERENCE = os.environ.get("B200TRAIN_FORCE_REFERENCE", "0") == "1"


def use_kernels(t: torch.Tensor) -> bool:
    if not t.is_cuda:
        return False
    if _FORCE_REFERENCE:
        return False
    _C.load()  # raises KernelError when the extension is missing on a GPU box
    return True


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream(t: torch.Tensor) -> int:
    """cudaStream_t of the current stream on ``t``'s device. The raw accessor skips building a
    ``torch.cuda.Stream`` object per kernel call (this function runs once per launch)."""
    if _raw_stream is not None:
        return _raw_stream(t.device.index if t.device.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(t.device).cuda_stream


def check_bf16(*tensors):
    for t in tensors:
        if t is not None and t.dtype != torch.bfloat16:
            raise TypeError(f"b200train kernels take bf16 tensors, got {t.dtype}")


def keep_params(ctx, **params):
    """Remember the parameter OBJECTS an autograd function was called with. ``ctx.saved_tensors`` does not
    always give them back: under non-reentrant activation checkpointing a saved parameter is replaced by a
    detached alias made during recomputation, which shares the storage but has neither ``.grad`` nor the
    direct-gradient flags (``_b2_direct_grad`` / ``_b2_grad_ready``) -- a backward that looked at the alias
    would hand the gradient to autograd and the unit's own bookkeeping would then zero it as "not produced"."""
    ctx._b2_params = params


def real_param(ctx, name, saved):
    p = getattr(ctx, "_b2_params", {}).get(name)
    return p if p is not None else saved
