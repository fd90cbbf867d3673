"""Flat parameter / gradient storage.

Every "unit" of a model (embedding, one decoder layer, final norm + head) keeps its parameters as views
into one contiguous bf16 buffer and its gradients as views into a second one. That gives:
  * wgrad GEMMs and norm backward kernels that write gradients in place (no ``.grad`` allocation,
    no zero-fill: ``zero_grad`` just clears the per-parameter "ready" flags),
  * one AdamW launch per unit,
  * one bucket per unit for the data-parallel gradient all-reduce, and one shard per unit for FSDP
    (the role of FSDP1's FlatParameter: ``04-fully-sharded-data-parallel/train_llm.py:87-95``).
"""
from __future__ import annotations

from typing import Iterable, List

import torch
from torch import nn

ALIGN = 128  # elements; keeps every parameter view 256-byte aligned for TMA and 16B vector access


def _round_up(n, a):
    return (n + a - 1) // a * a


class FlatUnit:
    """Flat storage for a list of parameters (deduplicated by identity, so tied weights appear once)."""

    def __init__(self, name: str, params: Iterable[nn.Parameter], pad_to: int = 1, with_grad: bool = True,
                 alloc=None):
        self.name = name
        self.params: List[nn.Parameter] = []
        seen = set()
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        assert self.params, f"unit {name} has no parameters"
        ref = self.params[0]
        self.device, self.dtype = ref.device, ref.dtype
        self.offsets = []
        off = 0
        for p in self.params:
            assert p.dtype == self.dtype and p.device == self.device
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.numel = _round_up(off, ALIGN * max(1, pad_to))
        # ``alloc(numel)`` lets a communicator place the buffers in symmetric (peer-mapped) memory
        alloc = alloc or (lambda k: torch.empty(k, dtype=self.dtype, device=self.device))
        self.flat_param = alloc(self.numel).zero_()
        self.flat_grad = alloc(self.numel).zero_() if with_grad else None
        # direct: kernels write gradients in place and flag them; otherwise autograd accumulates into
        # the flat views and zero_grad has to clear the buffer (CPU / reference-math path)
        self.direct = self.device.type == "cuda" and self.dtype == torch.bfloat16
        for p, o in zip(self.params, self.offsets):
            view = self.flat_param[o:o + p.numel()].view(p.shape)
            if p.device.type != "meta":
                view.copy_(p.data)
            p.data = view
            if with_grad:
                p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)
                p._b2_direct_grad = self.direct
                p._b2_grad_ready = False
            p._b2_unit = self

    def repoint(self, flat_param: torch.Tensor):
        """Make the parameters views of another buffer of the same layout (FSDP gathered copies)."""
        for p, o in zip(self.params, self.offsets):
            p.data = flat_param[o:o + p.numel()].view(p.shape)

    def zero_grad(self):
        """Gradients are overwritten by the next backward; only the ready flags are cleared."""
        if not self.direct:
            self.flat_grad.zero_()
            return
        for p in self.params:
            p._b2_grad_ready = False

    def finalize_grads(self):
        """Zero the gradient of any parameter the backward pass did not reach."""
        if not self.direct:
            return
        for p in self.params:
            if not p._b2_grad_ready:
                p.grad.zero_()

    def named_slices(self):
        for p, o in zip(self.params, self.offsets):
            yield p, o, p.numel()
