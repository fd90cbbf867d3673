"""Autograd boundaries around model units.

``run_unit`` brackets a unit (embedding, decoder layer, head) with two identity autograd functions:
the one on the unit's outputs fires first in backward ("pre-backward": parameters must be present,
gradient storage assigned) and the one on its inputs fires last ("post-backward": every gradient of the
unit has been written, so it can be reduced / the optimizer can run). This is the role of FSDP's
pre/post-backward hooks and of DDP's autograd-hook driven reducer in the reference
(``04-fully-sharded-data-parallel/train_llm.py:87-95``, ``02-distributed-data-parallel/train_llm.py:65``).
"""
from __future__ import annotations

import torch
from torch.utils.checkpoint import checkpoint

from ..obs import nvtx


class _PostBackward(torch.autograd.Function):
    """Identity on the unit inputs; its backward runs after the whole unit's backward."""

    @staticmethod
    def forward(ctx, owner, idx, *tensors):
        ctx.owner, ctx.idx = owner, idx
        return tuple(t.view_as(t) for t in tensors)

    @staticmethod
    def backward(ctx, *grads):
        nvtx.pop()  # closes unit<idx>.bwd opened by _PreBackward
        with nvtx.range(f"unit{ctx.idx}.post_bwd"):
            ctx.owner.post_backward(ctx.idx)
        return (None, None) + grads


class _PreBackward(torch.autograd.Function):
    """Identity on the unit outputs; its backward runs before the unit's backward. ``paired`` says a
    _PostBackward exists for the unit (it closes the NVTX range opened here)."""

    @staticmethod
    def forward(ctx, owner, idx, paired, *tensors):
        ctx.owner, ctx.idx, ctx.paired = owner, idx, paired
        return tuple(t.view_as(t) for t in tensors)

    @staticmethod
    def backward(ctx, *grads):
        ctx.owner.pre_backward(ctx.idx)
        if ctx.paired:
            nvtx.push(f"unit{ctx.idx}.bwd")
        return (None, None, None) + grads


def run_unit(owner, idx, fn, *tensors, activation_checkpoint: bool = False):
    """Execute ``fn(*tensors)`` as unit ``idx`` of ``owner`` with forward/backward callbacks.

    ``owner`` provides ``pre_forward(idx)``, ``post_forward(idx)``, ``pre_backward(idx)`` and
    ``post_backward(idx)``. Units without differentiable inputs (the embedding) get their
    post-backward from the strategy once ``loss.backward()`` returns.
    """
    owner.pre_forward(idx)
    grad_inputs = [t for t in tensors if torch.is_tensor(t) and t.requires_grad]
    track = torch.is_grad_enabled()
    if track and grad_inputs:
        wrapped = list(_PostBackward.apply(owner, idx, *grad_inputs))
        it = iter(wrapped)
        tensors = tuple(next(it) if (torch.is_tensor(t) and t.requires_grad) else t for t in tensors)
    nvtx.push(f"unit{idx}.fwd")
    if activation_checkpoint and track and grad_inputs:
        # reference: apply_activation_checkpointing on every decoder layer,
        # 05-training-llama-405b/train_llm.py:124-126 (recompute the layer forward in backward)
        out = checkpoint(fn, *tensors, use_reentrant=False)
    else:
        out = fn(*tensors)
    nvtx.pop()
    owner.post_forward(idx)
    if not track:
        return out
    single = torch.is_tensor(out)
    outs = (out,) if single else tuple(out)
    outs = _PreBackward.apply(owner, idx, bool(grad_inputs), *outs)
    return outs[0] if single else outs
