"""Linear layers on the tcgen05 GEMM (csrc/gemm.cu).

forward  y  = x W^T (+ bias)          K-major x K-major
dgrad    dx = dy W                    K-major x MN-major (no transposed copy of W)
wgrad    dW = dy^T x                  MN-major x MN-major, written straight into the parameter's
                                      flat gradient buffer when the parameter owns one

Reference: every ``nn.Linear`` (Llama) / ``Conv1D`` (GPT-2) reached from
``01-single-gpu/train_llm.py:140,145`` through HF modeling code and cuBLAS.
"""
from __future__ import annotations

import torch

from .. import _C
from .dispatch import check_bf16, keep_params, real_param, stream, use_kernels

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_ACCUM = 0, 1, 2, 3

# 0 = library default; (cta_group | BN << 8) to pin a tile configuration (used by benches)
GEMM_CONFIG = 0


class GatherWait:
    """Arrival flags of the flat unit buffer a weight lives in while FSDP is still gathering it
    (``parallel/fsdp.py``). ``offset`` is the weight's element offset in that buffer."""

    __slots__ = ("flags_ptr", "epoch", "offset", "chunk_elems", "n_chunks", "chunks_per_rank", "rank", "world")

    def __init__(self, flags_ptr, epoch, offset, chunk_elems, n_chunks, chunks_per_rank, rank, world):
        self.flags_ptr, self.epoch, self.offset, self.chunk_elems = flags_ptr, epoch, offset, chunk_elems
        self.n_chunks, self.chunks_per_rank, self.rank, self.world = n_chunks, chunks_per_rank, rank, world

    def stream_wait(self, numel: int, stream_ptr: int):
        """Order ``stream`` behind the arrival of every chunk that holds ``numel`` elements from ``offset``."""
        lo = self.offset // self.chunk_elems
        hi = min((self.offset + numel - 1) // self.chunk_elems, self.n_chunks - 1)
        _C.call("b2_stream_wait_flags", self.flags_ptr, lo, hi - lo + 1, self.epoch, stream_ptr)


def gemm_gathered(a: torch.Tensor, w: torch.Tensor, ag: GatherWait, *, b_mn=False, bias=None) -> torch.Tensor:
    """``a @ w^T`` (or ``a @ w`` with ``b_mn``) where ``w`` is still being all-gathered: the GEMM itself waits,
    tile by tile, for the chunks of the gathered buffer that hold the weight rows it is about to load
    (``b2_gemm_bf16_bwait``). Shapes the fused kernel does not cover wait on the stream and run the plain GEMM."""
    M = a.shape[0]
    N = w.shape[1] if b_mn else w.shape[0]
    K = a.shape[1]
    if M <= 128 or (b_mn and N % 64) or a.stride(1) != 1 or w.stride(1) != 1:
        ag.stream_wait(w.numel(), stream(a))
        return gemm(a, w, b_mn=b_mn, epilogue=EPI_BIAS if bias is not None else EPI_NONE, bias=bias)
    out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    _C.call("b2_gemm_bf16_bwait", out.data_ptr(), a.data_ptr(), w.data_ptr(), M, N, K, int(b_mn), a.stride(0),
            w.stride(0), out.stride(0), EPI_BIAS if bias is not None else EPI_NONE, _C.ptr(bias), ag.flags_ptr,
            ag.epoch, ag.offset, ag.chunk_elems, ag.n_chunks, ag.chunks_per_rank, ag.rank, ag.world, stream(a))
    return out


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn=False, b_mn=False, out=None, epilogue=EPI_NONE,
         bias=None, config=None, max_ctas=0) -> torch.Tensor:
    """C[M,N] = epilogue(A B^T). ``a``: [M,K] (or [K,M] when a_mn), ``b``: [N,K] (or [K,N] when b_mn).

    Operands are 2-D row-major views (unit stride in the last dim, arbitrary row stride)."""
    check_bf16(a, b, out, bias)
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"gemm: inner dims differ ({K} vs {Kb})"
    if (a_mn and M % 64) or (b_mn and N % 64):
        return _gemm_padded(a, b, a_mn, b_mn, out, epilogue, bias, config, max_ctas, M, N)
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    assert out.shape == (M, N) and out.stride(1) == 1
    _C.call("b2_gemm_bf16", out.data_ptr(), a.data_ptr(), b.data_ptr(), M, N, K, int(a_mn), int(b_mn),
            a.stride(0), b.stride(0), out.stride(0), epilogue, _C.ptr(bias),
            GEMM_CONFIG if config is None else config, max_ctas, stream(a))
    return out


def _gemm_padded(a, b, a_mn, b_mn, out, epilogue, bias, config, max_ctas, M, N):
    """MN-major operands are fetched in 64-element atoms: pad their contiguous extent to a multiple of 64
    with zeros (e.g. a vocabulary shard of 16032 columns under tp = 8), run, and keep the valid block."""
    assert epilogue in (EPI_NONE, EPI_ACCUM) and bias is None
    Mp, Np = (M + 63) // 64 * 64, (N + 63) // 64 * 64
    if a_mn and Mp != M:
        ap = torch.zeros((a.shape[0], Mp), dtype=a.dtype, device=a.device)
        ap[:, :M] = a
        a = ap
    else:
        Mp = M
    if b_mn and Np != N:
        bp = torch.zeros((b.shape[0], Np), dtype=b.dtype, device=b.device)
        bp[:, :N] = b
        b = bp
    else:
        Np = N
    tmp = gemm(a, b, a_mn=a_mn, b_mn=b_mn, config=config, max_ctas=max_ctas)[:M, :N]
    if out is None:
        return tmp.contiguous()
    if epilogue == EPI_ACCUM:
        out.add_(tmp)
    else:
        out.copy_(tmp)
    return out


def _direct_grad(w) -> bool:
    return getattr(w, "_b2_direct_grad", False) and w.grad is not None


_colsum_scratch = {}


def _colsum(x2: torch.Tensor, out: torch.Tensor, accumulate: bool):
    """out[N] (+)= x2.sum(0) with fp32 accumulation (``b2_colsum``)."""
    T, N = x2.shape
    lib = _C.load()
    need = lib.b2_colsum_partial_rows(T, N) * N
    buf = _colsum_scratch.get(x2.device)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.float32, device=x2.device)
        _colsum_scratch[x2.device] = buf
    _C.call("b2_colsum", out.data_ptr(), buf.data_ptr(), x2.data_ptr(), T, N, x2.stride(0), int(accumulate),
            stream(x2))


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, w_is_kn):
        # w_is_kn: weight stored [in, out] (GPT-2 Conv1D) instead of [out, in]
        # x is 2-D [T, in]; the output is a fresh (non-view) tensor so callers may modify it in place
        ctx.w_is_kn = w_is_kn
        ctx.save_for_backward(x, w, bias)
        keep_params(ctx, w=w, bias=bias)
        ag = getattr(w, "_b2_ag", None)
        if ag is not None:   # FSDP: the weight's all-gather may still be in flight
            return gemm_gathered(x, w, ag, b_mn=w_is_kn, bias=bias)
        return gemm(x, w, b_mn=w_is_kn, epilogue=EPI_BIAS if bias is not None else EPI_NONE, bias=bias)

    @staticmethod
    def backward(ctx, dy):
        x2, w, bias = ctx.saved_tensors
        w, bias = real_param(ctx, "w", w), real_param(ctx, "bias", bias)
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dx[M, in] = dy[M, out] * W ; W as the "B" operand is [in, out]-logical
            ag = getattr(w, "_b2_ag", None)
            dx = (gemm_gathered(dy2, w, ag, b_mn=not ctx.w_is_kn) if ag is not None
                  else gemm(dy2, w, b_mn=not ctx.w_is_kn))
        if ctx.needs_input_grad[1]:
            if ctx.w_is_kn:   # dW[in, out] = x^T dy
                a, b = x2, dy2
            else:             # dW[out, in] = dy^T x
                a, b = dy2, x2
            if _direct_grad(w):
                acc = getattr(w, "_b2_grad_ready", False)
                gemm(a, b, a_mn=True, b_mn=True, out=w.grad, epilogue=EPI_ACCUM if acc else EPI_NONE)
                w._b2_grad_ready = True
            else:
                dw = gemm(a, b, a_mn=True, b_mn=True)
        if bias is not None and ctx.needs_input_grad[2]:
            direct = _direct_grad(bias)
            acc = direct and getattr(bias, "_b2_grad_ready", False)
            db_out = bias.grad if direct else torch.empty_like(bias)
            _colsum(dy2, db_out, acc)      # one pass + fold, written straight into the flat gradient
            if direct:
                bias._b2_grad_ready = True
            else:
                db = db_out
        return dx, dw, db, None


def linear_ref(x, w, bias=None, w_is_kn=False):
    if w_is_kn:
        if bias is not None and x.dim() >= 2:
            return torch.addmm(bias, x.reshape(-1, x.shape[-1]), w).view(*x.shape[:-1], w.shape[1])
        y = x @ w
        return y + bias if bias is not None else y
    return torch.nn.functional.linear(x, w, bias)


def linear(x: torch.Tensor, w: torch.Tensor, bias=None, w_is_kn: bool = False) -> torch.Tensor:
    """y = x W^T + b for ``w`` [out, in]; with ``w_is_kn`` the weight is [in, out] (Conv1D layout)."""
    if use_kernels(x):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = _Linear.apply(x2, w, bias, w_is_kn)
        return y if x.dim() == 2 else y.view(*x.shape[:-1], y.shape[-1])
    return linear_ref(x, w, bias, w_is_kn)
