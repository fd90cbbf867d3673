"""SwiGLU on a fused gate|up activation and GPT-2 ``gelu_new`` (csrc/elementwise.cu).

Reference: HF ``LlamaMLP.forward`` ``down(act_fn(gate(x)) * up(x))`` and GPT-2 ``NewGELUActivation``.
"""
from __future__ import annotations


import torch

from .. import _C
from .dispatch import check_bf16, stream, use_kernels


class _SwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        check_bf16(gu)
        I = gu.shape[-1] // 2
        g2 = gu.reshape(-1, 2 * I)
        out = torch.empty((g2.shape[0], I), dtype=gu.dtype, device=gu.device)
        _C.call("b2_swiglu_fwd", out.data_ptr(), g2.data_ptr(), g2.shape[0], I, stream(gu))
        ctx.save_for_backward(g2)
        ctx.shape = gu.shape
        return out.view(*gu.shape[:-1], I)

    @staticmethod
    def backward(ctx, dout):
        (g2,) = ctx.saved_tensors
        I = g2.shape[1] // 2
        d2 = dout.reshape(-1, I)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        dgu = torch.empty_like(g2)
        _C.call("b2_swiglu_bwd", dgu.data_ptr(), d2.data_ptr(), g2.data_ptr(), g2.shape[0], I, stream(g2))
        return dgu.view(ctx.shape)


def swiglu_ref(gu):
    g, u = gu.chunk(2, dim=-1)
    return torch.nn.functional.silu(g) * u


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    """silu(gu[..., :I]) * gu[..., I:] for a fused gate|up projection output."""
    if use_kernels(gu):
        return _SwiGLU.apply(gu.contiguous())
    return swiglu_ref(gu)


class _GeluNew(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        check_bf16(x)
        x = x.contiguous()
        out = torch.empty_like(x)
        _C.call("b2_gelu_fwd", out.data_ptr(), x.data_ptr(), x.numel(), stream(x))
        ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, dout):
        (x,) = ctx.saved_tensors
        dout = dout.contiguous()
        dx = torch.empty_like(x)
        _C.call("b2_gelu_bwd", dx.data_ptr(), x.data_ptr(), dout.data_ptr(), x.numel(), stream(x))
        return dx


def gelu_new_ref(x):
    # HF "gelu_new" is the tanh approximation; ATen implements exactly that formula as one fused kernel
    return torch.nn.functional.gelu(x, approximate="tanh")


def gelu_new(x: torch.Tensor) -> torch.Tensor:
    if use_kernels(x) and x.numel() % 8 == 0:
        return _GeluNew.apply(x)
    return gelu_new_ref(x)
