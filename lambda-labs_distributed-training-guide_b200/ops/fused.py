"""GEMMs with their elementwise neighbours fused into the epilogue (csrc/gemm.cu).

``mlp_swiglu(x, w_gate_up, w_down)`` -- the whole Llama MLP as one autograd node:
  forward   GU = x W_gu^T and H = silu(gate) * up from ONE GEMM (``b2_gemm_swiglu``: the two CTAs of a pair
            load the gate and the up rows of the stacked weight, the epilogue writes both pre-activations
            and their SwiGLU), then y = H W_down^T;
  backward  d(gate|up) straight from the down-projection dgrad (``b2_gemm_dswiglu``: dh = dy W_down is
            combined with the saved pre-activations in the epilogue and never written), then the usual
            dgrad / wgrad GEMMs.
``linear_rope(x, w_qkv, ...)`` -- QKV projection with the rotary embedding of the q / k heads applied in the
  GEMM epilogue (``b2_gemm_rope``); backward un-rotates the incoming gradient in place and runs the
  ordinary dgrad / wgrad.

The standalone kernels (``ops.swiglu``, ``ops.rope_qk``) remain for shapes these do not cover and as the
oracle: the epilogues round the projection output to bf16 before the elementwise math, exactly as a
separate kernel reading the stored bf16 tensor would, so fused and unfused forwards agree bit for bit.

Reference: HF ``LlamaMLP.forward`` (``down(act_fn(gate(x)) * up(x))``) and ``apply_rotary_pos_emb`` behind
``01-single-gpu/train_llm.py:140``; the Liger-patched variants of ``00-rime/train_llm_01-single-gpu.py:51-54``.
"""
from __future__ import annotations

import os

import torch

from .. import _C
from .activation import swiglu
from .dispatch import check_bf16, keep_params, real_param, stream, use_kernels
from .linear import EPI_ACCUM, EPI_NONE, _direct_grad, gemm, gemm_gathered, linear
from .rope import rope_qk

ENABLED = os.environ.get("B200TRAIN_FUSED_EPILOGUES", "1") != "0"


def _gw(w):
    """ctypes arguments describing the in-flight all-gather of ``w`` (all zero / null when there is none)."""
    ag = getattr(w, "_b2_ag", None)
    if ag is None:
        return (None, 0, 0, 0, 0, 0, 0, 0)
    return (ag.flags_ptr, ag.epoch, ag.offset, ag.chunk_elems, ag.n_chunks, ag.chunks_per_rank, ag.rank, ag.world)


def _wgrad(a, b, w):
    """dW = a^T b, written into the parameter's flat gradient when it owns one (else returned)."""
    if _direct_grad(w):
        acc = getattr(w, "_b2_grad_ready", False)
        gemm(a, b, a_mn=True, b_mn=True, out=w.grad, epilogue=EPI_ACCUM if acc else EPI_NONE)
        w._b2_grad_ready = True
        return None
    return gemm(a, b, a_mn=True, b_mn=True)


def _dgrad(dy, w):
    ag = getattr(w, "_b2_ag", None)
    return gemm_gathered(dy, w, ag, b_mn=True) if ag is not None else gemm(dy, w, b_mn=True)


class _MLPSwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_gu, w_down):
        check_bf16(x, w_gu, w_down)
        T, H = x.shape
        I = w_gu.shape[0] // 2
        gu = torch.empty((T, 2 * I), dtype=torch.bfloat16, device=x.device)
        h = torch.empty((T, I), dtype=torch.bfloat16, device=x.device)
        _C.call("b2_gemm_swiglu", gu.data_ptr(), h.data_ptr(), x.data_ptr(), w_gu.data_ptr(), T, I, H, x.stride(0),
                w_gu.stride(0), gu.stride(0), h.stride(0), *_gw(w_gu), stream(x))
        ag = getattr(w_down, "_b2_ag", None)
        y = gemm_gathered(h, w_down, ag) if ag is not None else gemm(h, w_down)
        ctx.save_for_backward(x, gu, h, w_gu, w_down)
        keep_params(ctx, w_gu=w_gu, w_down=w_down)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gu, h, w_gu, w_down = ctx.saved_tensors
        w_gu, w_down = real_param(ctx, "w_gu", w_gu), real_param(ctx, "w_down", w_down)
        dy = dy.reshape(-1, dy.shape[-1])
        if not dy.is_contiguous():
            dy = dy.contiguous()
        T, H = x.shape
        I = h.shape[1]
        # d(gate|up) = dswiglu(dy W_down, gate|up): W_down [H, I] is the MN-major operand of its own dgrad
        dgu = torch.empty_like(gu)
        _C.call("b2_gemm_dswiglu", dgu.data_ptr(), dy.data_ptr(), w_down.data_ptr(), gu.data_ptr(), T, I, H,
                dy.stride(0), w_down.stride(0), dgu.stride(0), gu.stride(0), *_gw(w_down), stream(dy))
        dw_down = _wgrad(dy, h, w_down) if ctx.needs_input_grad[2] else None
        dx = _dgrad(dgu, w_gu) if ctx.needs_input_grad[0] else None
        dw_gu = _wgrad(dgu, x, w_gu) if ctx.needs_input_grad[1] else None
        return dx, dw_gu, dw_down


def mlp_swiglu(x: torch.Tensor, w_gate_up: torch.Tensor, w_down: torch.Tensor) -> torch.Tensor:
    """``down(silu(gate(x)) * up(x))`` for a stacked gate|up weight [2I, H] and a down weight [H, I]."""
    I = w_gate_up.shape[0] // 2
    x2 = x.reshape(-1, x.shape[-1])
    if ENABLED and use_kernels(x) and x2.shape[0] > 128 and I % 128 == 0 and x2.shape[1] % 64 == 0:
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = _MLPSwiGLU.apply(x2, w_gate_up, w_down)
        return y if x.dim() == 2 else y.view(*x.shape[:-1], y.shape[-1])
    return linear(swiglu(linear(x, w_gate_up)), w_down)


class _LinearRope(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, cos, sin, pos, nheads, seq_len):
        check_bf16(x, w, cos, sin)
        T, K = x.shape
        N = w.shape[0]
        out = torch.empty((T, N), dtype=torch.bfloat16, device=x.device)
        _C.call("b2_gemm_rope", out.data_ptr(), x.data_ptr(), w.data_ptr(), T, N, K, x.stride(0), w.stride(0),
                out.stride(0), cos.data_ptr(), sin.data_ptr(), _C.ptr(pos), nheads, seq_len, *_gw(w), stream(x))
        ctx.save_for_backward(x, w, cos, sin, pos)
        keep_params(ctx, w=w)
        ctx.args = (nheads, seq_len)
        return out

    @staticmethod
    def backward(ctx, dqkv):
        x, w, cos, sin, pos = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        nheads, seq_len = ctx.args
        dqkv = dqkv.contiguous()
        # transpose of the rotation, in place on the incoming gradient (its only consumer is this node)
        _C.call("b2_rope", dqkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), _C.ptr(pos), dqkv.shape[0], nheads, 128,
                dqkv.shape[1], seq_len, 1, stream(dqkv))
        dx = _dgrad(dqkv, w) if ctx.needs_input_grad[0] else None
        dw = _wgrad(dqkv, x, w) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None, None, None, None


def linear_rope(x, w, cos, sin, pos, nheads: int, head_dim: int, seq_len: int) -> torch.Tensor:
    """``rope_qk(linear(x, w))``: fused QKV projection with the first ``nheads`` heads rotated."""
    x2 = x.reshape(-1, x.shape[-1])
    if (ENABLED and use_kernels(x) and head_dim == 128 and x2.shape[0] > 128 and w.shape[0] % 128 == 0
            and cos.shape[1] == 128):
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        return _LinearRope.apply(x2, w, cos, sin, pos, nheads, seq_len)
    return rope_qk(linear(x, w), cos, sin, pos, nheads, head_dim, seq_len)
