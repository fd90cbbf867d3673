"""LM-head projection + shifted-label cross entropy without fp32 logits.

The head GEMM writes bf16 logits (padded to a multiple of 64 columns so the same buffer can be the
MN-major operand of the head's wgrad GEMM); one kernel computes the per-row log-sum-exp / loss and
one kernel turns the logits into dlogits in place. The reference materialises fp32 logits
(``logits.float()`` in HF ``ForCausalLMLoss``, 2.1 GB at T=4096, V=128256) and runs ATen
log_softmax + nll_loss (``01-single-gpu/train_llm.py:140`` -> ``outputs.loss``).

``vocab_parallel_cross_entropy`` is the tensor-parallel variant (reference:
``06-tensor-parallel/train_llm.py:117-127,239-243`` ``loss_parallel``): each rank holds a vocabulary
shard of the logits; the per-row max / sum-exp / target-logit statistics are exchanged over the tp
group instead of gathering logits.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from .. import _C
from .dispatch import check_bf16, keep_params, real_param, stream, use_kernels
from .linear import gemm, EPI_ACCUM, EPI_NONE, _direct_grad

IGNORE = -100


def shift_labels(labels: torch.Tensor) -> torch.Tensor:
    """HF convention: position t predicts token t+1; the last position of each row is ignored."""
    shifted = torch.full_like(labels, IGNORE)
    shifted[..., :-1] = labels[..., 1:]
    return shifted


class _LinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, labels):
        """x [T,H] bf16, w [V,H] bf16, labels [T] int64 already shifted. Returns mean loss (fp32)."""
        check_bf16(x, w)
        T, H = x.shape
        V = w.shape[0]
        ld = (V + 63) // 64 * 64
        logits = torch.empty((T, ld), dtype=torch.bfloat16, device=x.device)
        gemm(x, w, out=logits[:, :V])
        lse = torch.empty(T, dtype=torch.float32, device=x.device)
        loss_rows = torch.empty(T, dtype=torch.float32, device=x.device)
        labels = labels.contiguous()
        _C.call("b2_ce_fwd", lse.data_ptr(), loss_rows.data_ptr(), logits.data_ptr(), labels.data_ptr(),
                T, V, ld, IGNORE, stream(x))
        n_valid = (labels != IGNORE).sum().clamp_(min=1)
        loss = loss_rows.sum() / n_valid
        ctx.save_for_backward(x, w, labels, logits, lse, n_valid)
        keep_params(ctx, w=w)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        x, w, labels, logits, lse, n_valid = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        T, H = x.shape
        V = w.shape[0]
        ld = logits.shape[1]
        # dlogits in place; the scale dloss / n_valid stays on the device (no host sync)
        scale_t = (dloss.float() / n_valid.float()).reshape(1).contiguous()
        _C.call("b2_ce_bwd", logits.data_ptr(), logits.data_ptr(), lse.data_ptr(), labels.data_ptr(), T, V,
                ld, IGNORE, 0, 1.0, scale_t.data_ptr(), stream(x))
        dl = logits[:, :V]
        dx = gemm(dl, w, b_mn=True) if ctx.needs_input_grad[0] else None
        dw = None
        if ctx.needs_input_grad[1]:
            # dW[V,H] = dlogits^T x ; the padded columns of dlogits are zero
            a = logits
            if _direct_grad(w):
                acc = getattr(w, "_b2_grad_ready", False)
                out = w.grad
                if ld != V:
                    tmp = gemm(a, x, a_mn=True, b_mn=True)
                    if acc:
                        out.add_(tmp[:V])
                    else:
                        out.copy_(tmp[:V])
                else:
                    gemm(a, x, a_mn=True, b_mn=True, out=out, epilogue=EPI_ACCUM if acc else EPI_NONE)
                w._b2_grad_ready = True
            else:
                dw = gemm(a, x, a_mn=True, b_mn=True)[:V]
        return dx, dw, None


def ce_chunk_rows_default() -> int:
    """Rows of logits alive at once in ``linear_cross_entropy`` (0: all of them, the fastest). Set by
    ``--ce-chunk-rows`` / ``B200TRAIN_CE_CHUNK_ROWS``."""
    return int(os.environ.get("B200TRAIN_CE_CHUNK_ROWS", "0"))


class _ChunkedLinearCrossEntropy(torch.autograd.Function):
    """LM head + cross entropy that never holds more than ``chunk`` rows of logits.

    Forward walks the token rows in chunks: head GEMM into ONE reused ``[chunk, V]`` bf16 buffer, then the
    row kernel leaves only the log-sum-exp and the loss of every row (8 bytes per token). Backward walks the
    same chunks: the logits of a chunk are recomputed into the same buffer, turned into dlogits in place, and
    consumed at once by the two gradient GEMMs (``dx`` rows of the chunk; ``dW`` accumulated in place through
    the GEMM's accumulate epilogue, straight into the flat gradient buffer when the strategy owns one). Cost:
    one extra head GEMM (2 T V H FLOP) and one read + write of dW per chunk -- measured 9.0 -> 15.0 ms at T=4096,
    52.1 -> 65.3 ms at T=16384 (V=128256, chunk 2048; ``profiles/ce_chunk_bench_r2.jsonl``); saved:
    ``(T - chunk) * V * 2`` bytes of activations (0.98 -> 0.49 GB and 3.91 -> 0.49 GB), which is what the reference's
    fork buys with Liger's fused linear cross entropy (``00-rime/train_llm_01-single-gpu.py:51-54``). The same
    code runs on CPU tensors with ATen pieces, which is how the chunk bookkeeping is unit-tested.
    """

    @staticmethod
    def _logits(xc, w, buf, V):
        if use_kernels(xc):
            gemm(xc, w, out=buf[:xc.shape[0], :V])
        else:
            torch.mm(xc, w.t(), out=buf[:xc.shape[0], :V])
        return buf[:xc.shape[0]]

    @staticmethod
    def forward(ctx, x, w, labels, chunk):
        T, H = x.shape
        V = w.shape[0]
        ld = (V + 63) // 64 * 64
        kernels = use_kernels(x)
        if kernels:
            check_bf16(x, w)
        chunk = max(1, min(int(chunk), T))
        labels = labels.contiguous()
        buf = torch.empty((chunk, ld), dtype=x.dtype, device=x.device)
        lse = torch.empty(T, dtype=torch.float32, device=x.device)
        loss_rows = torch.empty(T, dtype=torch.float32, device=x.device)
        for r0 in range(0, T, chunk):
            r1 = min(T, r0 + chunk)
            lg = _ChunkedLinearCrossEntropy._logits(x[r0:r1], w, buf, V)
            if kernels:
                _C.call("b2_ce_fwd", lse[r0:r1].data_ptr(), loss_rows[r0:r1].data_ptr(), lg.data_ptr(),
                        labels[r0:r1].data_ptr(), r1 - r0, V, ld, IGNORE, stream(x))
            else:
                lf = lg[:, :V].float()
                lse[r0:r1] = torch.logsumexp(lf, dim=1)
                lab = labels[r0:r1]
                tgt = lf.gather(1, lab.clamp(min=0)[:, None])[:, 0]
                loss_rows[r0:r1] = torch.where(lab != IGNORE, lse[r0:r1] - tgt, torch.zeros_like(tgt))
        n_valid = (labels != IGNORE).sum().clamp_(min=1)
        ctx.save_for_backward(x, w, labels, lse, n_valid)
        ctx.chunk, ctx.ld = chunk, ld
        keep_params(ctx, w=w)
        del buf
        return loss_rows.sum() / n_valid

    @staticmethod
    def backward(ctx, dloss):
        x, w, labels, lse, n_valid = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        T, H = x.shape
        V = w.shape[0]
        chunk, ld = ctx.chunk, ctx.ld
        kernels = use_kernels(x)
        scale_t = (dloss.float() / n_valid.float()).reshape(1).contiguous()
        buf = torch.empty((chunk, ld), dtype=x.dtype, device=x.device)
        need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dx = torch.empty_like(x) if need_dx else None
        direct = need_dw and kernels and _direct_grad(w)
        # accumulator of dW: the strategy's flat gradient when it owns one, else a fresh tensor
        if direct:
            dw_out, started = w.grad, bool(getattr(w, "_b2_grad_ready", False))
        elif need_dw:
            dw_out, started = torch.empty((ld if kernels else V, H), dtype=x.dtype, device=x.device), False
        for r0 in range(0, T, chunk):
            r1 = min(T, r0 + chunk)
            xc = x[r0:r1]
            lg = _ChunkedLinearCrossEntropy._logits(xc, w, buf, V)
            if kernels:
                _C.call("b2_ce_bwd", lg.data_ptr(), lg.data_ptr(), lse[r0:r1].data_ptr(),
                        labels[r0:r1].data_ptr(), r1 - r0, V, ld, IGNORE, 0, 1.0, scale_t.data_ptr(), stream(x))
            else:
                lab = labels[r0:r1]
                p = torch.exp(lg[:, :V].float() - lse[r0:r1, None])
                rows = (lab != IGNORE).nonzero()[:, 0]
                p[rows, lab[rows]] -= 1.0
                p[lab == IGNORE] = 0
                lg[:, :V] = (p * scale_t).to(lg.dtype)
            dl = lg[:, :V]
            if need_dx:
                if kernels:
                    gemm(dl, w, b_mn=True, out=dx[r0:r1])
                else:
                    torch.mm(dl, w, out=dx[r0:r1])
            if need_dw:
                if not kernels:
                    dw_out.copy_(dl.t() @ xc) if not started else dw_out.add_(dl.t() @ xc)
                elif direct and ld != V:
                    tmp = gemm(lg, xc, a_mn=True, b_mn=True)   # [ld, H]: the padded rows are zero
                    dw_out.add_(tmp[:V]) if started else dw_out.copy_(tmp[:V])
                else:
                    gemm(lg, xc, a_mn=True, b_mn=True, out=dw_out, epilogue=EPI_ACCUM if started else EPI_NONE)
                started = True
        dw = None
        if direct:
            w._b2_grad_ready = True
        elif need_dw:
            dw = dw_out[:V]
        return dx, dw, None, None


def linear_cross_entropy_ref(x, w, labels):
    logits = torch.nn.functional.linear(x, w).float()
    return torch.nn.functional.cross_entropy(logits, labels, ignore_index=IGNORE)


def linear_cross_entropy(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, chunk_rows=None) -> torch.Tensor:
    """mean CE of ``x @ w.T`` against already-shifted ``labels`` (ignore_index=-100). ``chunk_rows`` > 0 (default:
    ``B200TRAIN_CE_CHUNK_ROWS``) keeps at most that many rows of logits alive (recomputed in backward)."""
    x2 = x.reshape(-1, x.shape[-1])
    l1 = labels.reshape(-1)
    chunk = ce_chunk_rows_default() if chunk_rows is None else int(chunk_rows)
    if 0 < chunk < x2.shape[0]:
        return _ChunkedLinearCrossEntropy.apply(x2.contiguous(), w, l1, chunk)
    if use_kernels(x2):
        return _LinearCrossEntropy.apply(x2.contiguous(), w, l1)
    return linear_cross_entropy_ref(x2, w, l1)


# ------------------------------------------------------------------------------------------------
# vocabulary-parallel variant
# ------------------------------------------------------------------------------------------------
class _VocabParallelCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, vocab_start, group, peer=None):
        """logits [T, Vlocal] (this rank's vocabulary shard), labels [T] global ids (shifted). ``peer``: the
        tp group's peer-memory handle (``parallel/tp_fused.FusedTP``); with it the shard statistics are
        combined by a kernel that reads the peers' memory instead of two NCCL all-reduces."""
        T, Vl = logits.shape
        labels = labels.contiguous()
        if use_kernels(logits):
            check_bf16(logits)
            assert logits.is_contiguous() and Vl % 8 == 0
            stats = torch.empty((3, T), dtype=torch.float32, device=logits.device)
            _C.call("b2_ce_shard_stats", stats[0].data_ptr(), stats[1].data_ptr(), stats[2].data_ptr(),
                    logits.data_ptr(), labels.data_ptr(), T, Vl, Vl, IGNORE, vocab_start, stream(logits))
            lmax, lsum, ltgt = stats[0], stats[1], stats[2]
        else:
            lf = logits.float()
            lmax = lf.max(dim=1).values
            lsum = torch.exp(lf - lmax[:, None]).sum(1)
            local = labels - vocab_start
            mine = (labels != IGNORE) & (local >= 0) & (local < Vl)
            ltgt = torch.where(mine, lf.gather(1, local.clamp(0, Vl - 1)[:, None])[:, 0], torch.zeros_like(lmax))
        valid = labels != IGNORE
        n_valid = valid.sum().clamp_(min=1)
        if peer is not None and use_kernels(logits):
            lse, loss_rows = peer.ce_combine(stats, labels)
            loss = loss_rows.sum() / n_valid
        else:
            gmax = lmax.clone()
            dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=group)
            packed = torch.stack((lsum * torch.exp(lmax - gmax), ltgt))
            dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
            lse = gmax + torch.log(packed[0])
            loss = torch.where(valid, lse - packed[1], torch.zeros_like(lse)).sum() / n_valid
        ctx.save_for_backward(logits, labels, lse, n_valid)
        ctx.vocab_start = vocab_start
        return loss

    @staticmethod
    def backward(ctx, dloss):
        logits, labels, lse, n_valid = ctx.saved_tensors
        T, Vl = logits.shape
        scale_t = (dloss.float() / n_valid.float()).reshape(1).contiguous()
        if use_kernels(logits):
            dl = torch.empty_like(logits)
            _C.call("b2_ce_bwd", dl.data_ptr(), logits.data_ptr(), lse.data_ptr(), labels.data_ptr(), T, Vl,
                    Vl, IGNORE, ctx.vocab_start, 1.0, scale_t.data_ptr(), stream(logits))
        else:
            p = torch.exp(logits.float() - lse[:, None])
            local = labels - ctx.vocab_start
            mine = (labels != IGNORE) & (local >= 0) & (local < Vl)
            p[mine, local[mine]] -= 1.0
            p[labels == IGNORE] = 0
            dl = (p * scale_t).to(logits.dtype)
        return dl, None, None, None, None


def vocab_parallel_cross_entropy(logits, labels, vocab_start: int, group=None, peer=None):
    """Mean CE over vocabulary-sharded ``logits`` [T, V/tp]; ``labels`` are global, already shifted."""
    return _VocabParallelCE.apply(logits, labels.reshape(-1), vocab_start, group, peer)
