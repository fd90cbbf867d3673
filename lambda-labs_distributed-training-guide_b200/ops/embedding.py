"""Token embedding gather and its scatter-add backward (csrc/elementwise.cu).

Reference: ``nn.Embedding`` inside HF ``LlamaModel`` / ``GPT2Model`` (``01-single-gpu/train_llm.py:140``).
"""
from __future__ import annotations

import torch

from .. import _C
from .dispatch import keep_params, real_param, check_bf16, stream, use_kernels


class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, table):
        check_bf16(table)
        flat = ids.reshape(-1).to(torch.int64).contiguous()
        H = table.shape[1]
        out = torch.empty((flat.numel(), H), dtype=table.dtype, device=table.device)
        _C.call("b2_embedding_fwd", out.data_ptr(), table.data_ptr(), flat.data_ptr(), flat.numel(), H,
                stream(table))
        ctx.save_for_backward(flat, table)
        keep_params(ctx, table=table)
        return out.view(*ids.shape, H)

    @staticmethod
    def backward(ctx, dout):
        flat, table = ctx.saved_tensors
        table = real_param(ctx, "table", table)
        H = table.shape[1]
        d2 = dout.reshape(-1, H)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        direct = getattr(table, "_b2_direct_grad", False) and table.grad is not None
        accumulate = 0
        if direct:
            if getattr(table, "_b2_grad_ready", False):
                accumulate = 1          # e.g. a tied lm_head already wrote its wgrad into this buffer
            else:
                table.grad.zero_()
            table._b2_grad_ready = True
            dt = table.grad
        else:
            dt = torch.zeros_like(table)
        # sort the ids once: runs of equal ids are reduced in fp32 by one block each (no atomics)
        sorted_ids, perm = torch.sort(flat)
        _C.call("b2_embedding_bwd", dt.data_ptr(), d2.data_ptr(), sorted_ids.data_ptr(), perm.data_ptr(),
                flat.numel(), H, accumulate, stream(table))
        return None, (None if direct else dt)


def embedding(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    if use_kernels(table):
        return _Embedding.apply(ids, table)
    return torch.nn.functional.embedding(ids, table)
