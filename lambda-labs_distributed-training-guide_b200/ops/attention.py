"""Causal flash attention over a fused QKV activation (csrc/attention_fwd.cu, attention_bwd.cu).

``attention(qkv, n_q, n_kv, head_dim, cu_seqlens, max_seqlen)`` takes the post-RoPE projection
output [T, (Hq + 2 Hkv) * D] and returns [T, Hq * D]; q/k/v are strided views of that buffer and the
backward writes dq/dk/dv into one fused gradient buffer, so the QKV GEMM's dgrad consumes it as is.
Packed batches are described by ``cu_seqlens`` (int32 [ndocs+1]).

Reference: HF attention dispatch to ``flash_attn`` / SDPA (``05-training-llama-405b/train_llm.py:87``,
``01-single-gpu/train_llm.py:140``); varlen via ``position_ids`` restarts (``00-rime/packed_dataset.py:35-48``).
"""
from __future__ import annotations

import math
import os

import torch

from .. import _C
from .dispatch import check_bf16, stream, use_kernels

# "kernel": tcgen05 backward (csrc/attention_bwd.cu); "flash_attn": library backward (debug only)
BWD_IMPL = os.environ.get("B200TRAIN_ATTN_BWD", "kernel")


def _split(qkv, n_q, n_kv, D):
    q = qkv[:, : n_q * D]
    k = qkv[:, n_q * D: (n_q + n_kv) * D]
    v = qkv[:, (n_q + n_kv) * D:]
    return q, k, v


def attn_fwd_raw(q, k, v, cu_seqlens, max_seqlen, n_q, n_kv, D, scale, causal=True, dropout=None):
    """Forward on strided [T, heads*D] views; returns (o [T, Hq*D], lse [Hq, T]). ``dropout``: None or
    ``(p, seed, offset)`` for dropout on the attention probabilities."""
    T = q.shape[0]
    o = torch.empty((T, n_q * D), dtype=torch.bfloat16, device=q.device)
    lse = torch.empty((n_q, T), dtype=torch.float32, device=q.device)
    if dropout is not None:
        _C.call("b2_attn_fwd_dropout", o.data_ptr(), lse.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(),
                cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_seqlen, T, n_q, n_kv, D, q.stride(0),
                k.stride(0), v.stride(0), o.stride(0), scale, int(causal), *dropout, stream(q))
        return o, lse
    _C.call("b2_attn_fwd", o.data_ptr(), lse.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(),
            cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_seqlen, T, n_q, n_kv, D, q.stride(0),
            k.stride(0), v.stride(0), o.stride(0), scale, int(causal), stream(q))
    return o, lse


def attn_bwd_raw(do, q, k, v, o, lse, dq, dk, dv, cu_seqlens, max_seqlen, n_q, n_kv, D, scale, causal=True,
                 dropout=None):
    T = q.shape[0]
    if dropout is not None:
        dq_accum = torch.empty((T, n_q * D), dtype=torch.float32, device=q.device)
        delta = torch.empty((n_q, T), dtype=torch.float32, device=q.device)
        _C.call("b2_attn_bwd_dropout", dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dq_accum.data_ptr(),
                delta.data_ptr(), do.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(),
                lse.data_ptr(), cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_seqlen, T, n_q, n_kv, D,
                q.stride(0), k.stride(0), v.stride(0), o.stride(0), do.stride(0), dq.stride(0), dk.stride(0),
                dv.stride(0), scale, int(causal), *dropout, stream(q))
        return
    if BWD_IMPL == "flash_attn":
        from flash_attn.flash_attn_interface import _flash_attn_varlen_backward

        _flash_attn_varlen_backward(
            do.view(T, n_q, D), q.unflatten(1, (n_q, D)), k.unflatten(1, (n_kv, D)), v.unflatten(1, (n_kv, D)),
            o.view(T, n_q, D), lse, dq.unflatten(1, (n_q, D)), dk.unflatten(1, (n_kv, D)),
            dv.unflatten(1, (n_kv, D)), cu_seqlens, cu_seqlens, max_seqlen, max_seqlen, 0.0, scale, causal,
            -1, -1, 0.0, None, False, None)
        return
    dq_accum = torch.empty((T, n_q * D), dtype=torch.float32, device=q.device)
    delta = torch.empty((n_q, T), dtype=torch.float32, device=q.device)
    _C.call("b2_attn_bwd", dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dq_accum.data_ptr(), delta.data_ptr(),
            do.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(),
            cu_seqlens.data_ptr(), cu_seqlens.numel() - 1, max_seqlen, T, n_q, n_kv, D, q.stride(0), k.stride(0),
            v.stride(0), o.stride(0), do.stride(0), dq.stride(0), dk.stride(0), dv.stride(0), scale, int(causal),
            stream(q))


class _Attention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cu_seqlens, max_seqlen, n_q, n_kv, D, scale, causal, dropout=None):
        check_bf16(qkv)
        assert qkv.dim() == 2 and qkv.is_contiguous()
        q, k, v = _split(qkv, n_q, n_kv, D)
        o, lse = attn_fwd_raw(q, k, v, cu_seqlens, max_seqlen, n_q, n_kv, D, scale, causal, dropout)
        ctx.save_for_backward(qkv, o, lse, cu_seqlens)
        ctx.args = (max_seqlen, n_q, n_kv, D, scale, causal, dropout)
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse, cu_seqlens = ctx.saved_tensors
        max_seqlen, n_q, n_kv, D, scale, causal, dropout = ctx.args
        do = do.contiguous()
        dqkv = torch.empty_like(qkv)
        q, k, v = _split(qkv, n_q, n_kv, D)
        dq, dk, dv = _split(dqkv, n_q, n_kv, D)
        attn_bwd_raw(do, q, k, v, o, lse, dq, dk, dv, cu_seqlens, max_seqlen, n_q, n_kv, D, scale, causal, dropout)
        return dqkv, None, None, None, None, None, None, None, None


_MASKS = {}


def _causal_mask(L: int) -> torch.Tensor:
    """[L, L] bool, True above the diagonal (cached per length: the CPU path builds it once, not per layer)."""
    m = _MASKS.get(L)
    if m is None:
        if len(_MASKS) > 8:
            _MASKS.clear()
        m = _MASKS[L] = torch.ones(L, L, dtype=torch.bool).triu_(1)
    return m


def attention_ref(qkv, cu_seqlens, max_seqlen, n_q, n_kv, D, scale=None, causal=True, dropout_p=0.0):
    """fp32 PyTorch reference over each document separately (``dropout_p``: torch's own mask)."""
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    T = qkv.shape[0]
    q, k, v = _split(qkv, n_q, n_kv, D)
    q = q.reshape(T, n_q, D)
    k = k.reshape(T, n_kv, D)
    v = v.reshape(T, n_kv, D)
    outs = []
    bounds = cu_seqlens.tolist()
    g = n_q // n_kv
    for s, e in zip(bounds[:-1], bounds[1:]):
        if qkv.device.type == "cpu" and qkv.dtype in (torch.float32, torch.bfloat16) and torch.is_grad_enabled():
            # the CPU plumbing path (chapter 01 on CPU): ATen's fused attention, no S x S tensor
            qi = q[s:e].transpose(0, 1).unsqueeze(0)
            ki = k[s:e].transpose(0, 1).unsqueeze(0)
            vi = v[s:e].transpose(0, 1).unsqueeze(0)
            if g > 1:
                ki, vi = ki.repeat_interleave(g, dim=1), vi.repeat_interleave(g, dim=1)
            if dropout_p > 0.0:
                # ATen's fused CPU attention has no dropout and its math fallback draws the mask from one
                # sequential stream (half of a GPT-2 CPU step): explicit products + the Philox dropout of ops.dropout
                from .dropout import dropout as _dropout

                att = torch.matmul(qi * scale, ki.transpose(-1, -2))   # scale on [L, D], not on [L, L]
                if causal:
                    att.masked_fill_(_causal_mask(e - s), float("-inf"))
                pi = _dropout(torch.softmax(att, dim=-1), dropout_p)
                oi = torch.matmul(pi, vi)
            else:
                oi = torch.nn.functional.scaled_dot_product_attention(qi, ki, vi, is_causal=causal, scale=scale)
            outs.append(oi[0].transpose(0, 1).reshape(e - s, n_q * D))
            continue
        qi = q[s:e].transpose(0, 1).float()
        ki = k[s:e].transpose(0, 1).repeat_interleave(g, dim=0).float()
        vi = v[s:e].transpose(0, 1).repeat_interleave(g, dim=0).float()
        att = (qi @ ki.transpose(-1, -2)) * scale
        if causal:
            L = e - s
            mask = torch.ones(L, L, dtype=torch.bool, device=qkv.device).tril()
            att = att.masked_fill(~mask, float("-inf"))
        pi = torch.softmax(att, dim=-1)
        if dropout_p > 0.0:
            pi = torch.nn.functional.dropout(pi, dropout_p)
        oi = pi @ vi
        outs.append(oi.transpose(0, 1).reshape(e - s, n_q * D))
    return torch.cat(outs, 0).to(qkv.dtype)


def attention(qkv: torch.Tensor, cu_seqlens: torch.Tensor, max_seqlen: int, n_q: int, n_kv: int,
              head_dim: int, scale: float | None = None, causal: bool = True, dropout_p: float = 0.0,
              philox=None) -> torch.Tensor:
    """``dropout_p`` > 0: dropout on the attention probabilities inside the kernels (forward applies a Philox
    mask keyed on (seed, offset, head, query, key); backward regenerates it). ``philox``: explicit
    ``(seed, offset)``, else drawn from the default CPU generator."""
    scale = scale if scale is not None else 1.0 / math.sqrt(head_dim)
    if use_kernels(qkv):
        drop = None
        if dropout_p > 0.0:
            from .dropout import next_philox

            seed, offset = philox if philox is not None else next_philox()
            drop = (float(dropout_p), seed, offset)
        return _Attention.apply(qkv, cu_seqlens, max_seqlen, n_q, n_kv, head_dim, scale, causal, drop)
    return attention_ref(qkv, cu_seqlens, max_seqlen, n_q, n_kv, head_dim, scale, causal, dropout_p)
