"""Dropout with a counter-based (Philox) mask that backward regenerates instead of storing.

GPT-2 is the only model family of the reference with dropout: HF's ``GPT2Config`` defaults to 0.1 at three
sites (embedding sum, attention probabilities, the two residual branches) and ``from_config`` leaves the module
in training mode (``02-distributed-data-parallel/train_llm.py:60-62``). The elementwise sites use
``csrc/elementwise.cu:dropout_kernel``; the attention-probability site lives inside the attention kernels
(``ops.attention(..., dropout_p=...)``).

The (seed, offset) pair of a call comes from PyTorch's CPU generator, so ``torch.manual_seed`` makes runs
repeatable and activation checkpointing (which restores the generator before recomputing) replays the same
masks.
"""
from __future__ import annotations

import torch

from .. import _C
from .dispatch import check_bf16, stream, use_kernels

_MASK63 = (1 << 63) - 1


def next_philox():
    """A fresh (seed, offset) for one dropout site: two draws from the default CPU generator."""
    t = torch.empty(2, dtype=torch.int64).random_()
    return int(t[0]) & _MASK63, int(t[1]) & _MASK63


class _Dropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed, offset):
        check_bf16(x)
        x = x.contiguous()
        out = torch.empty_like(x)
        _C.call("b2_dropout", out.data_ptr(), x.data_ptr(), x.numel(), p, seed, offset, stream(x))
        ctx.args = (p, seed, offset)
        return out

    @staticmethod
    def backward(ctx, dout):
        p, seed, offset = ctx.args
        dout = dout.contiguous()
        dx = torch.empty_like(dout)
        _C.call("b2_dropout", dx.data_ptr(), dout.data_ptr(), dout.numel(), p, seed, offset, stream(dout))
        return dx, None, None, None


def dropout(x: torch.Tensor, p: float, training: bool = True) -> torch.Tensor:
    if p <= 0.0 or not training:
        return x
    if use_kernels(x) and x.numel() % 8 == 0:
        seed, offset = next_philox()
        return _Dropout.apply(x, float(p), seed, offset)
    return torch.nn.functional.dropout(x, p, training=True)
