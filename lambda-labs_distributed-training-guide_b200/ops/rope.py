"""Rotary position embedding applied in place to the q and k heads of a fused QKV activation
(csrc/elementwise.cu ``rope_kernel``), plus the Llama-3 scaled inverse-frequency tables.

Reference: HF ``apply_rotary_pos_emb`` / ``LlamaRotaryEmbedding`` (llama3 rope scaling) reached from
``01-single-gpu/train_llm.py:140``; explicit position ids as in ``06-tensor-parallel/train_llm.py:231-233``
and the packed collator ``00-rime/packed_dataset.py:35``.
"""
from __future__ import annotations

import math

import torch

from .. import _C
from .dispatch import check_bf16, stream, use_kernels


def rope_tables(head_dim: int, max_pos: int, theta: float = 10000.0, scaling: dict | None = None,
                device=None, dtype=torch.bfloat16):
    """cos/sin tables [max_pos, head_dim] (HF layout: the half-dim frequencies repeated twice)."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    if scaling and scaling.get("rope_type", scaling.get("type")) == "llama3":
        factor = scaling["factor"]
        low, high = scaling["low_freq_factor"], scaling["high_freq_factor"]
        old_len = scaling["original_max_position_embeddings"]
        wavelen = 2 * math.pi / inv_freq
        scaled = torch.where(wavelen > old_len / low, inv_freq / factor, inv_freq)
        smooth = (old_len / wavelen - low) / (high - low)
        smoothed = (1 - smooth) * scaled / factor + smooth * scaled
        mid = ~(wavelen < old_len / high) & ~(wavelen > old_len / low)
        inv_freq = torch.where(mid, smoothed, scaled)
    t = torch.arange(max_pos, dtype=torch.float32)
    freqs = torch.outer(t, inv_freq)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype=dtype, device=device), emb.sin().to(dtype=dtype, device=device)


class _RopeQK(torch.autograd.Function):
    """In-place rotation of the first ``nheads`` heads of each row of ``qkv`` [T, row]."""

    @staticmethod
    def forward(ctx, qkv, cos, sin, pos, nheads, head_dim, seq_len):
        check_bf16(qkv, cos, sin)
        assert qkv.is_contiguous() and qkv.dim() == 2
        _C.call("b2_rope", qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), _C.ptr(pos), qkv.shape[0],
                nheads, head_dim, qkv.shape[1], seq_len, 0, stream(qkv))
        ctx.save_for_backward(cos, sin, pos)
        ctx.args = (nheads, head_dim, seq_len)
        ctx.mark_dirty(qkv)
        return qkv

    @staticmethod
    def backward(ctx, dqkv):
        cos, sin, pos = ctx.saved_tensors
        nheads, head_dim, seq_len = ctx.args
        dqkv = dqkv.contiguous()
        _C.call("b2_rope", dqkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), _C.ptr(pos), dqkv.shape[0],
                nheads, head_dim, dqkv.shape[1], seq_len, 1, stream(dqkv))
        return dqkv, None, None, None, None, None, None


def _rotate_half(x):
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def rope_qk_ref(qkv, cos, sin, pos, nheads, head_dim, seq_len):
    T = qkv.shape[0]
    if pos is None:
        pos = torch.arange(T, device=qkv.device) % seq_len
    c = cos[pos.long()].unsqueeze(1).float()
    s = sin[pos.long()].unsqueeze(1).float()
    x = qkv[:, : nheads * head_dim].reshape(T, nheads, head_dim).float()
    rot = (x * c + _rotate_half(x) * s).to(qkv.dtype).reshape(T, nheads * head_dim)
    return torch.cat((rot, qkv[:, nheads * head_dim:]), dim=1)


def rope_qk(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos, nheads: int, head_dim: int,
            seq_len: int) -> torch.Tensor:
    """Rotate heads [0, nheads) (q heads then k heads) of ``qkv`` [T, (Hq+2Hkv)*D]; v is untouched.

    ``pos``: int32 [T] positions (packed sequences) or None for ``t % seq_len``."""
    if use_kernels(qkv):
        return _RopeQK.apply(qkv, cos, sin, pos, nheads, head_dim, seq_len)
    return rope_qk_ref(qkv, cos, sin, pos, nheads, head_dim, seq_len)
