"""Fused tensor-parallel linears over NVLink peer memory (csrc/gemm.cu ``b2_gemm_ag`` / ``b2_gemm_rs``).

``colwise(x_local, w)``  = all-gather(sequence) -> GEMM. The local rows are written into a symmetric
    buffer, the copy engines pull the peers' chunks on a side stream (one completion flag per chunk),
    and ONE tcgen05 GEMM kernel starts at once: it computes the tiles of the local chunk first and its
    TMA producer waits on a chunk's flag before touching it, so transfer and math overlap tile by tile.
``rowwise(x, w)``        = GEMM -> reduce-scatter(sequence). The GEMM epilogue TMA-stores every output
    tile directly into the owner rank's staging slot (peer-mapped memory) while the main loop continues;
    after a device-side barrier each rank sums its W slots.
Backward mirrors them (dX of colwise is a rowwise-style GEMM->RS, dX of rowwise an AG->GEMM) and the
weight gradients reuse the gathered operand.

Reference: ``ColwiseParallel`` / ``RowwiseParallel(output_layouts=Shard(1))`` with
``PrepareModuleInput(Shard(1) -> Replicate())`` (``06-tensor-parallel/train_llm.py:89-115``), which run
NCCL all-gather / reduce-scatter followed by (or following) a cuBLAS GEMM.
"""
from __future__ import annotations

import torch

from .. import _C
from ..ops.dispatch import keep_params, real_param
from ..ops.linear import EPI_ACCUM, EPI_NONE, _direct_grad, gemm
from .peer import PeerComm

CHAN_AG_BARRIER, CHAN_RS_BARRIER, CHAN_SMALL_BARRIER = 6, 7, 10


class FusedTP:
    """Buffers, flags and streams shared by the fused TP linears of one tp group."""

    def __init__(self, tp_group, rank: int, size: int, device):
        self.rank, self.W, self.device = rank, size, torch.device(device)
        self.comm = PeerComm(tp_group, self.device)
        self.comm_stream = torch.cuda.Stream(self.device)
        self.flags = torch.zeros(16, dtype=torch.int32, device=self.device)
        self.ag_epoch = 0
        self._ag_bufs = {}    # (rows, cols) -> [buf0, buf1]
        self._slot_bufs = {}  # (rows_per_rank, cols) -> [buf0, buf1]
        self._ag_i = 0
        self._rs_i = 0
        self._bar_epochs = {CHAN_AG_BARRIER: 0, CHAN_RS_BARRIER: 0, CHAN_SMALL_BARRIER: 0}
        self._small = {}      # name -> [symmetric buf0, buf1] (double buffered: a rank may run one call ahead)
        self._small_i = {}

    # ---- buffers (symmetric, two alternating copies; see module docstring of csrc/gemm.cu) ----------
    def ag_buffer(self, rows: int, cols: int) -> torch.Tensor:
        key = (rows, cols)
        if key not in self._ag_bufs:
            self._ag_bufs[key] = [self.comm.empty(rows * cols).view(rows, cols) for _ in range(2)]
        self._ag_i ^= 1
        return self._ag_bufs[key][self._ag_i]

    def slot_buffer(self, rows_per_rank: int, cols: int) -> torch.Tensor:
        key = (rows_per_rank, cols)
        if key not in self._slot_bufs:
            self._slot_bufs[key] = [self.comm.empty(self.W * rows_per_rank * cols).view(self.W, rows_per_rank, cols)
                                    for _ in range(2)]
        self._rs_i ^= 1
        return self._slot_bufs[key][self._rs_i]

    def _barrier(self, chan: int):
        self._bar_epochs[chan] += 1
        _C.call("b2_peer_barrier", self.comm._pad_ptrs, self.rank, self.W, chan, self._bar_epochs[chan] & 0xFFFFFFFF,
                torch.cuda.current_stream(self.device).cuda_stream)

    # ---- the plan's small collectives, in device code / on the copy engines (N9, N10, N11) -------------------
    def _small_buf(self, name: str, numel: int, dtype) -> torch.Tensor:
        """Symmetric staging buffer ``name`` of at least ``numel`` elements; alternates between two copies so
        a rank that is one call ahead never overwrites what a slower peer is still reading (the barrier in
        every call bounds the lead to one)."""
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        pair = self._small.get(name)
        if pair is None or pair[0].numel() < nbytes:
            # every rank reaches this with the same sizes in the same order (symmetric allocation)
            pair = [self.comm.empty(nbytes, torch.uint8) for _ in range(2)]
            self._small[name] = pair
            self._small_i[name] = 0
        self._small_i[name] ^= 1
        return pair[self._small_i[name]][:nbytes].view(dtype)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def all_reduce_sum_(self, tensors):
        """Sum a list of small bf16 tensors over the tp group in ONE in-switch (or peer) all-reduce: they are
        packed into a symmetric buffer, reduced, and copied back (N11: the sequence-parallel norm-weight
        gradients of a unit, or of the whole model)."""
        sizes = [t.numel() for t in tensors]
        total = sum(sizes)
        q = 8 * self.W
        padded = (total + q - 1) // q * q
        buf = self._small_buf("ar", padded, torch.bfloat16)
        if padded != total:
            buf[total:].zero_()
        views = list(buf[:total].split(sizes))
        torch._foreach_copy_(views, [t.reshape(-1) for t in tensors])
        self.comm.all_reduce(buf, average=False)
        for t, v in zip(tensors, views):
            t.copy_(v.view(t.shape))

    def ce_combine(self, stats: torch.Tensor, labels: torch.Tensor):
        """``stats`` fp32 [3, T] (row max, sum exp(x - max), target logit of this rank's vocabulary shard) ->
        (lse [T], loss_row [T]) over the whole vocabulary, each rank reading the peers' statistics from peer
        memory inside one kernel (N10)."""
        T = stats.shape[1]
        buf = self._small_buf("ce", 3 * T, torch.float32)
        buf.copy_(stats.reshape(-1))
        self._barrier(CHAN_SMALL_BARRIER)
        lse = torch.empty(T, dtype=torch.float32, device=self.device)
        loss_row = torch.empty(T, dtype=torch.float32, device=self.device)
        _C.call("b2_ce_combine_peer", lse.data_ptr(), loss_row.data_ptr(), self.comm._peer_array(buf),
                labels.data_ptr(), T, self.W, -100, self._stream())
        return lse, loss_row

    def hidden_to_seq(self, x: torch.Tensor) -> torch.Tensor:
        """[T, H/tp] (all tokens, local hidden slice) -> [T/tp, H] (local tokens, full hidden): every rank
        pulls its token rows of every peer's slice with strided copy-engine copies (N9)."""
        T, Hl = x.shape
        n = T // self.W
        buf = self._small_buf("h2s", T * Hl, x.dtype)
        buf.copy_(x.reshape(-1))
        self._barrier(CHAN_SMALL_BARRIER)
        out = torch.empty((n, self.W * Hl), dtype=x.dtype, device=self.device)
        es = x.element_size()
        _C.call("b2_peer_pull_2d", out.data_ptr(), self.W * Hl * es, Hl * es, self.comm._peer_array(buf),
                self.rank * n * Hl * es, Hl * es, Hl * es, n, self.rank, self.W, self._stream())
        return out

    def seq_to_hidden(self, g: torch.Tensor) -> torch.Tensor:
        """Transpose of :meth:`hidden_to_seq` (its backward): [T/tp, H] -> [T, H/tp]."""
        n, H = g.shape
        Hl = H // self.W
        buf = self._small_buf("s2h", n * H, g.dtype)
        buf.copy_(g.reshape(-1))
        self._barrier(CHAN_SMALL_BARRIER)
        out = torch.empty((self.W * n, Hl), dtype=g.dtype, device=self.device)
        es = g.element_size()
        _C.call("b2_peer_pull_2d", out.data_ptr(), Hl * es, n * Hl * es, self.comm._peer_array(buf),
                self.rank * Hl * es, H * es, Hl * es, n, self.rank, self.W, self._stream())
        return out

    # ---- all-gather of row chunks with per-chunk flags ------------------------------------------------
    def gather_rows(self, x_local: torch.Tensor):
        """Returns (full buffer [W*n, cols], epoch, done_event). The local chunk is in place when this
        returns (stream order); peer chunk r is in place once ``flags[r] >= epoch``."""
        n, cols = x_local.shape
        buf = self.ag_buffer(n * self.W, cols)
        buf[self.rank * n:(self.rank + 1) * n].copy_(x_local)
        ready = torch.cuda.Event()
        ready.record()
        self.ag_epoch += 1
        bases = self.comm._peer_array(buf)
        with torch.cuda.stream(self.comm_stream):
            self.comm_stream.wait_event(ready)
            self._barrier(CHAN_AG_BARRIER)        # every rank's chunk is written
            after_barrier = torch.cuda.Event()
            after_barrier.record()
            _C.call("b2_peer_all_gather_chunks", buf.data_ptr(), bases, n * cols * 2, self.rank, self.W,
                    self.flags.data_ptr(), self.ag_epoch, self.comm_stream.cuda_stream)
            done = torch.cuda.Event()
            done.record()
        # the consumer GEMM may only depend on copy-engine work and stream memory ops (neither needs an
        # SM): it is released after the cross-rank barrier kernel has finished, never while it spins
        torch.cuda.current_stream(self.device).wait_event(after_barrier)
        return buf, self.ag_epoch, done

    def gemm_ag(self, buf, epoch, w, b_mn=False):
        M, K = buf.shape
        N = w.shape[1] if b_mn else w.shape[0]
        out = torch.empty((M, N), dtype=torch.bfloat16, device=self.device)
        _C.call("b2_gemm_ag", out.data_ptr(), buf.data_ptr(), w.data_ptr(), M, N, K, int(b_mn), buf.stride(0),
                w.stride(0), out.stride(0), self.flags.data_ptr(), epoch, M // self.W, self.rank, self.W,
                torch.cuda.current_stream(self.device).cuda_stream)
        return out

    def gemm_rs(self, x, w, b_mn=False):
        """RS(x @ w^T) -> [M/W, N]: scatter in the GEMM epilogue, then barrier + local slot sum."""
        M, K = x.shape
        N = w.shape[1] if b_mn else w.shape[0]
        n = M // self.W
        slots = self.slot_buffer(n, N)
        bases = self.comm._peer_array(slots)
        import ctypes

        mine = self.rank * n * N * 2   # slot `rank` inside every peer's staging buffer
        slot_ptrs = (ctypes.c_void_p * self.W)(*[bases[r] + mine for r in range(self.W)])
        _C.call("b2_gemm_rs", slot_ptrs, x.data_ptr(), w.data_ptr(), M, N, K, int(b_mn), x.stride(0), w.stride(0),
                N, n, self.rank, self.W, torch.cuda.current_stream(self.device).cuda_stream)
        self._barrier(CHAN_RS_BARRIER)            # every rank's tiles have landed in my slots
        out = torch.empty((n, N), dtype=torch.bfloat16, device=self.device)
        _C.call("b2_sum_slots", out.data_ptr(), slots.data_ptr(), n * N, self.W,
                torch.cuda.current_stream(self.device).cuda_stream)
        return out


def _wgrad(a, b, w):
    """dW = a^T b written into the parameter's flat gradient when it owns one."""
    if _direct_grad(w):
        acc = getattr(w, "_b2_grad_ready", False)
        gemm(a, b, a_mn=True, b_mn=True, out=w.grad, epilogue=EPI_ACCUM if acc else EPI_NONE)
        w._b2_grad_ready = True
        return None
    return gemm(a, b, a_mn=True, b_mn=True)


class _ColwiseFused(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_local, w, fz: FusedTP):
        x_local = x_local.contiguous()
        buf, epoch, _ = fz.gather_rows(x_local)
        y = fz.gemm_ag(buf, epoch, w)
        ctx.save_for_backward(x_local, w)
        keep_params(ctx, w=w)
        ctx.fz = fz
        return y

    @staticmethod
    def backward(ctx, dy):
        x_local, w = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        fz = ctx.fz
        dy = dy.contiguous()
        # the gathered activations are needed again for dW: start re-gathering, overlap with the dX GEMM
        buf, epoch, done = fz.gather_rows(x_local)
        dx_local = fz.gemm_rs(dy, w, b_mn=True)            # dX = RS(dy W)
        torch.cuda.current_stream(fz.device).wait_event(done)
        dw = _wgrad(dy, buf, w)                            # dW = dy^T AG(x)
        return dx_local, dw, None


class _RowwiseFused(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, fz: FusedTP):
        x = x.contiguous()
        y_local = fz.gemm_rs(x, w)
        ctx.save_for_backward(x, w)
        keep_params(ctx, w=w)
        ctx.fz = fz
        return y_local

    @staticmethod
    def backward(ctx, dy_local):
        x, w = ctx.saved_tensors
        w = real_param(ctx, "w", w)
        fz = ctx.fz
        buf, epoch, done = fz.gather_rows(dy_local.contiguous())
        dx = fz.gemm_ag(buf, epoch, w, b_mn=True)          # dX = AG(dy) W
        torch.cuda.current_stream(fz.device).wait_event(done)
        dw = _wgrad(buf, x, w)                             # dW = AG(dy)^T x
        return dx, dw, None


def colwise(x_local, w, tp):
    """y[T, N_local] = all_gather_seq(x_local) @ w^T."""
    fz = getattr(tp, "fused", None)
    if fz is not None and x_local.is_cuda and x_local.shape[0] % 256 == 0:
        return _ColwiseFused.apply(x_local, w, fz)
    from ..ops import linear
    from .tp_ops import all_gather_seq

    return linear(all_gather_seq(x_local, tp), w)


def rowwise(x, w, tp):
    """y_local[T/tp, N] = reduce_scatter_seq(x @ w^T)."""
    fz = getattr(tp, "fused", None)
    if fz is not None and x.is_cuda and (x.shape[0] // tp.size) % 256 == 0:
        return _RowwiseFused.apply(x, w, fz)
    from ..ops import linear
    from .tp_ops import reduce_scatter_seq

    return reduce_scatter_seq(linear(x, w), tp)
