"""RMSNorm / LayerNorm with an optional fused residual add (csrc/elementwise.cu).

``rmsnorm(x, w, eps, residual)`` returns ``(y, h)`` where ``h = x + residual`` (or ``x``) is the updated
residual stream, so one kernel covers "add the block output back, then normalise for the next block".
Reference math: HF ``LlamaRMSNorm.forward`` (fp32 variance, cast, then weight multiply) and GPT-2
``nn.LayerNorm`` reached from ``01-single-gpu/train_llm.py:140``.
"""
from __future__ import annotations

import torch

from .. import _C
from .dispatch import check_bf16, keep_params, real_param, stream, use_kernels

_scratch = {}


def _partial_scratch(device, nfloats):
    buf = _scratch.get(device)
    if buf is None or buf.numel() < nfloats:
        buf = torch.empty(nfloats, dtype=torch.float32, device=device)
        _scratch[device] = buf
    return buf


def _grad_target(w, new_shape_like):
    """(out tensor, accumulate flag, return-to-autograd?) for a small parameter gradient."""
    if getattr(w, "_b2_direct_grad", False) and w.grad is not None:
        acc = getattr(w, "_b2_grad_ready", False)
        w._b2_grad_ready = True
        return w.grad, acc, False
    return torch.empty_like(new_shape_like), False, True


class _RMSNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, residual, eps):
        check_bf16(x, w, residual)
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        T = x2.shape[0]
        r2 = residual.reshape(-1, H) if residual is not None else None
        y = torch.empty_like(x2)
        h = torch.empty_like(x2) if r2 is not None else x2
        rstd = torch.empty(T, dtype=torch.float32, device=x.device)
        _C.call("b2_rmsnorm_fwd", y.data_ptr(), rstd.data_ptr(), h.data_ptr() if r2 is not None else None,
                x2.data_ptr(), _C.ptr(r2), w.data_ptr(), T, H, eps, stream(x))
        ctx.save_for_backward(h, w, rstd)
        keep_params(ctx, w=w)
        ctx.has_res = residual is not None
        ctx.shape = x.shape
        if residual is None:
            return y.view(x.shape)
        return y.view(x.shape), h.view(x.shape)

    @staticmethod
    def backward(ctx, dy, dh=None):
        h, w, rstd = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        T, H = h.shape
        dy2 = dy.reshape(T, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dh2 = None
        if dh is not None:
            dh2 = dh.reshape(T, H)
            if not dh2.is_contiguous():
                dh2 = dh2.contiguous()
        dx = torch.empty_like(h)
        dw, acc, ret = _grad_target(w, w)
        nrows = _C.load().b2_rmsnorm_bwd_partial_rows(T)
        partial = _partial_scratch(h.device, nrows * H)
        _C.call("b2_rmsnorm_bwd", dx.data_ptr(), dw.data_ptr(), partial.data_ptr(), dy2.data_ptr(),
                h.data_ptr(), w.data_ptr(), rstd.data_ptr(), _C.ptr(dh2), T, H, int(acc), stream(h))
        dx = dx.view(ctx.shape)
        return dx, (dw if ret else None), (dx if ctx.has_res else None), None


def rmsnorm_ref(x, w, eps, residual=None):
    h = x + residual if residual is not None else x
    hf = h.float()
    var = hf.pow(2).mean(-1, keepdim=True)
    y = w * (hf * torch.rsqrt(var + eps)).to(x.dtype)
    return y, h


def rmsnorm(x, w, eps: float = 1e-5, residual=None):
    """Returns (normalised, residual_stream)."""
    if use_kernels(x):
        if residual is None:
            return _RMSNorm.apply(x, w, None, eps), x
        return _RMSNorm.apply(x, w, residual, eps)
    return rmsnorm_ref(x, w, eps, residual)


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, residual, eps):
        check_bf16(x, w, b, residual)
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        T = x2.shape[0]
        r2 = residual.reshape(-1, H) if residual is not None else None
        y = torch.empty_like(x2)
        h = torch.empty_like(x2) if r2 is not None else x2
        mean = torch.empty(T, dtype=torch.float32, device=x.device)
        rstd = torch.empty(T, dtype=torch.float32, device=x.device)
        _C.call("b2_layernorm_fwd", y.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                h.data_ptr() if r2 is not None else None, x2.data_ptr(), _C.ptr(r2), w.data_ptr(),
                b.data_ptr(), T, H, eps, stream(x))
        ctx.save_for_backward(h, w, b, mean, rstd)
        keep_params(ctx, w=w, b=b)
        ctx.has_res = residual is not None
        ctx.shape = x.shape
        if residual is None:
            return y.view(x.shape)
        return y.view(x.shape), h.view(x.shape)

    @staticmethod
    def backward(ctx, dy, dh=None):
        h, w, b, mean, rstd = ctx.saved_tensors
        w, b = real_param(ctx, "w", w), real_param(ctx, "b", b)
        T, H = h.shape
        dy2 = dy.reshape(T, H).contiguous()
        dh2 = dh.reshape(T, H).contiguous() if dh is not None else None
        dx = torch.empty_like(h)
        direct = getattr(w, "_b2_direct_grad", False) and w.grad is not None
        if direct:
            dw, db = w.grad, b.grad
            acc = getattr(w, "_b2_grad_ready", False)
            w._b2_grad_ready = True
            b._b2_grad_ready = True
        else:
            dw, db, acc = torch.empty_like(w), torch.empty_like(b), False
        nrows = _C.load().b2_rmsnorm_bwd_partial_rows(T)
        partial = _partial_scratch(h.device, nrows * 2 * H)
        _C.call("b2_layernorm_bwd", dx.data_ptr(), dw.data_ptr(), db.data_ptr(), partial.data_ptr(),
                dy2.data_ptr(), h.data_ptr(), w.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                _C.ptr(dh2), T, H, int(acc), stream(h))
        dx = dx.view(ctx.shape)
        return dx, (None if direct else dw), (None if direct else db), (dx if ctx.has_res else None), None


def layernorm_ref(x, w, b, eps, residual=None):
    h = x + residual if residual is not None else x
    y = torch.nn.functional.layer_norm(h.float(), (h.shape[-1],), w.float(), b.float(), eps).to(x.dtype)
    return y, h


def layernorm(x, w, b, eps: float = 1e-5, residual=None):
    """Returns (normalised, residual_stream)."""
    if use_kernels(x):
        if residual is None:
            return _LayerNorm.apply(x, w, b, None, eps), x
        return _LayerNorm.apply(x, w, b, residual, eps)
    return layernorm_ref(x, w, b, eps, residual)
