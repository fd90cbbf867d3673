"""Peer-memory (NVLink / NVSwitch) communicator for the ranks of one node.

Buffers that take part in collectives are carved out of *symmetric arenas*: CUDA allocations that every
rank of the group maps into its own address space (``torch.distributed._symmetric_memory`` does the VMM
allocation and the handle exchange -- plumbing only; every byte moved afterwards is moved by the
kernels in ``csrc/comm.cu`` with loads from the mapped peer pointers and flags in a symmetric signal pad).

``PeerComm`` implements the :class:`Comm` surface used by the DDP / FSDP strategies plus the fused
``reduce_scatter_adamw``. Tensors passed to its collectives must come from :meth:`empty` (strategies
allocate shards and gradient buffers through ``comm.empty``); anything else falls back to NCCL.

``B200TRAIN_COMM_DEBUG=1`` turns on a checking mode for the hand-written cross-GPU signalling (the
reference has nothing to check: SURVEY 5.2): every reduce-scatter / all-reduce / all-gather is repeated
through NCCL on copies of its inputs and the two results are compared on the device; a mismatch beyond
bf16 rounding raises with the collective's name, rank and worst element. It synchronises every call and
is meant for tests and bring-up, not for timing.

Reference: the implicit NCCL all-gather / reduce-scatter / all-reduce of FSDP and DDP
(``04-fully-sharded-data-parallel/train_llm.py:87-95``, ``02-distributed-data-parallel/train_llm.py:65``).
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _C
from ..obs import nvtx
from .comm import TorchComm

_ARENA_MIN_BYTES = 256 << 20
_ALIGN = 256
CHAN_RS, CHAN_AR, CHAN_BARRIER, CHAN_CE_IN, CHAN_CE_OUT = 0, 2, 4, 8, 9   # 6, 7, 10: tp_fused barriers
CHAN_AR_EXIT = 11


class _Arena:
    def __init__(self, nbytes: int, device, group):
        import torch.distributed._symmetric_memory as symm

        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=device)
        self.hdl = symm.rendezvous(self.buf, group=group if group is not None else dist.group.WORLD)
        self.base = self.buf.data_ptr()
        self.peer_bases = [int(p) for p in self.hdl.buffer_ptrs]
        # multicast (NVLS) mapping of the same allocation; 0 when the fabric has no multicast support
        self.mc_base = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        self.nbytes = nbytes
        self.used = 0

    def take(self, nbytes: int) -> Optional[torch.Tensor]:
        start = (self.used + _ALIGN - 1) // _ALIGN * _ALIGN
        if start + nbytes > self.nbytes:
            return None
        self.used = start + nbytes
        return self.buf[start:start + nbytes]


class PeerComm(TorchComm):
    def __init__(self, group=None, device=None):
        super().__init__(group)
        if self.backend != "nccl":
            raise RuntimeError("PeerComm needs CUDA ranks (nccl process group)")
        if self.world > 16:
            raise RuntimeError("peer-memory collectives support up to 16 ranks on one node")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.arenas: List[_Arena] = []
        # pure copies (parameter all-gather) can ride the DMA engines instead of SM load/store loops
        self.copy_engine = os.environ.get("B200TRAIN_AG", "ce") == "ce"
        # reductions through the NVSwitch (multimem.ld_reduce on the multicast mapping) when the arena has
        # one. "auto": from 4 ranks up, where one switch-side load replaces W-1 peer loads per element
        # (measured: profiles/comm_bench_w2.json, comm_bench_w4.json); at 2 ranks the peer loads are faster
        mode = os.environ.get("B200TRAIN_NVLS", "auto")
        self.nvls = mode == "1" or (mode == "auto" and self.world >= 4)
        # the switch saturates with few blocks in flight; fewer blocks leave more of the SMs to the GEMMs
        self.nvls_blocks = int(os.environ.get("B200TRAIN_NVLS_BLOCKS", "16"))
        self.rs_blocks = int(os.environ.get("B200TRAIN_RS_BLOCKS", "64"))
        # blocks of the fused reduce-scatter + AdamW kernel: wide at 2 ranks (the peer slice is half the unit and
        # the update is HBM bound), thin through the switch (each rank's NVLink egress bounds it anyway)
        self.rsadam_blocks = int(os.environ.get("B200TRAIN_RSADAM_BLOCKS", "0"))
        # the own shard of an all-gather is placed by this many copy blocks (0: cudaMemcpyAsync -- its copy kernel
        # holds back the GEMM launches of the compute stream while it runs, 75 us per 218 MB, but copy blocks
        # beside the GEMMs cost more: 200 vs 208-210 ms per Llama-8B step at 2 ranks)
        self.self_copy_blocks = int(os.environ.get("B200TRAIN_SELF_COPY_BLOCKS", "0"))
        self.rs_blocks = int(os.environ.get("B200TRAIN_RS_BLOCKS", "64"))
        # reduce-scatter transport. "ce": the copy engines pull the peers' slices into local staging and one
        # local kernel reduces (+ AdamW) -- nothing resident beside the GEMMs but an HBM-bound kernel; "nvls":
        # in-switch reduction by a thin kernel; "peer": SM loads from the peers. auto: see _transport
        self.rs_transport = os.environ.get("B200TRAIN_RS_TRANSPORT", "auto")
        self._ce_staging = {}
        self.debug = os.environ.get("B200TRAIN_COMM_DEBUG", "0") == "1"
        self.debug_checks = 0
        pad = self.empty(64 * 16, torch.int32)   # signal pad: [channel][16 ranks]
        pad.zero_()
        self._pad = pad
        self._pad_ptrs = self._peer_array(pad)
        self._epochs = {}
        self._counters = {}
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)          # pads are zeroed everywhere before the first signal

    # ---- symmetric allocation ---------------------------------------------------------------------
    def empty(self, numel: int, dtype=torch.bfloat16, device=None) -> torch.Tensor:
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        for a in self.arenas:
            t = a.take(nbytes)
            if t is not None:
                return t.view(dtype)
        # every rank allocates the same sequence of sizes, so arenas line up across ranks
        arena = _Arena(max(_ARENA_MIN_BYTES, nbytes + _ALIGN), self.device, self.group)
        self.arenas.append(arena)
        return arena.take(nbytes).view(dtype)

    def _locate(self, t: torch.Tensor):
        p = t.data_ptr()
        for a in self.arenas:
            if a.base <= p < a.base + a.nbytes:
                return a, p - a.base
        return None, 0

    def is_symmetric(self, t: torch.Tensor) -> bool:
        return self._locate(t)[0] is not None

    def _peer_array(self, t: torch.Tensor):
        a, off = self._locate(t)
        if a is None:
            return None
        arr = (ctypes.c_void_p * self.world)(*[b + off for b in a.peer_bases])
        return arr

    def _mc_ptr(self, t: torch.Tensor) -> int:
        """Multicast address of a symmetric tensor (0: not symmetric, or no NVLS on this machine)."""
        a, off = self._locate(t)
        return a.mc_base + off if a is not None and a.mc_base else 0

    @property
    def has_nvls(self) -> bool:
        return bool(self.arenas) and all(a.mc_base for a in self.arenas)

    def _epoch(self, chan: int) -> int:
        e = self._epochs.get(chan, 0) + 1
        self._epochs[chan] = e
        return e & 0xFFFFFFFF

    def _counter(self) -> torch.Tensor:
        key = torch.cuda.current_stream(self.device).cuda_stream
        c = self._counters.get(key)
        if c is None:
            c = torch.zeros(1, dtype=torch.int32, device=self.device)
            self._counters[key] = c
        return c

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    # ---- collectives --------------------------------------------------------------------------------
    def stream_barrier(self):
        if self.world == 1:
            return
        _C.call("b2_peer_barrier", self._pad_ptrs, self.rank, self.world, CHAN_BARRIER,
                self._epoch(CHAN_BARRIER), self._stream())

    def _all_gather(self, full, shard, max_blocks: int = 64, copy_engine: Optional[bool] = None):
        """``full`` is local; ``shard`` must be symmetric. Source shards must be stable: the caller
        orders them against their last writer with :meth:`barrier` (FSDP does this once per step)."""
        srcs = self._peer_array(shard) if self.world > 1 else None
        if srcs is None:
            return TorchComm.all_gather(self, full, shard)
        n = shard.numel() * shard.element_size()
        skip_self = int(full.data_ptr() + self.rank * n == shard.data_ptr())
        if self.copy_engine if copy_engine is None else copy_engine:
            _C.call("b2_peer_all_gather_ce", full.data_ptr(), srcs, n, self.rank, self.world, skip_self,
                    self.self_copy_blocks, self._stream())
            return
        _C.call("b2_peer_all_gather", full.data_ptr(), srcs, n, self.rank, self.world, skip_self, max_blocks,
                self._stream())

    def all_gather_flagged(self, full, shard, flags, epoch: int, sub: int, part: int = 2):
        """Copy-engine gather of ``shard`` (symmetric) into the local ``full`` with an arrival flag per chunk
        (``sub`` chunks per rank, ``flags[r * sub + j] = epoch`` written by the stream after the chunk).
        ``part``: 0 own shard only, 1 peers only, 2 both."""
        srcs = self._peer_array(shard)
        if srcs is None:
            raise RuntimeError("all_gather_flagged needs a symmetric shard (allocate it with comm.empty)")
        n = shard.numel() * shard.element_size()
        with nvtx.range("comm.all_gather_flagged"):
            _C.call("b2_peer_all_gather_flagged", full.data_ptr(), srcs, n, self.rank, self.world, sub,
                    self.self_copy_blocks, part, flags.data_ptr(), epoch, self._stream())

    def all_gather_inplace(self, full, max_blocks: int = 64):
        """In-place variant: slice q is pulled from rank q's copy of the same buffer (the owned slices sit
        at different offsets on different ranks, so the sources are ``peer_base[q] + q * slice``)."""
        bases = self._peer_array(full) if self.world > 1 else None
        if bases is None:
            return super().all_gather_inplace(full)
        n = full.numel() // self.world * full.element_size()
        srcs = (ctypes.c_void_p * self.world)(*[bases[q] + q * n for q in range(self.world)])
        if self.copy_engine:
            _C.call("b2_peer_all_gather_ce", full.data_ptr(), srcs, n, self.rank, self.world, 1, 0, self._stream())
        else:
            _C.call("b2_peer_all_gather", full.data_ptr(), srcs, n, self.rank, self.world, 1, max_blocks,
                    self._stream())

    # ---- copy-engine reduce-scatter ---------------------------------------------------------------------
    def _transport(self, nvls: Optional[bool], mc_available: bool, fused: bool = False) -> str:
        if nvls is not None:
            return "nvls" if (nvls and mc_available) else "peer"
        t = self.rs_transport
        if t == "auto":
            # fused with AdamW: copy-engine pull + local reduce-update at every world size (Llama-8B step on 4
            # GPUs: 198.4 ms, against 209.2 ms with the in-switch load feeding AdamW in one thin kernel and
            # 198.6 ms with in-switch reduce-scatter then a separate AdamW launch). Plain reduce-scatter: in-switch
            # from 4 ranks, copy engine at 2.
            if fused:
                t = "ce"
            else:
                t = "ce" if self.world == 2 else ("nvls" if (self.nvls and mc_available) else "ce")
        if t == "nvls" and not mc_available:
            t = "ce"
        return t

    def _ce_stage(self, nbytes: int) -> torch.Tensor:
        key = self._stream()
        buf = self._ce_staging.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            self._ce_staging[key] = buf
        return buf

    def _reduce_scatter_ce(self, fulls, shard_elems: int, local_slice: torch.Tensor, consume):
        """Handshake + copy-engine pull of the peers' copies of this rank's slice, then ``consume(staging)``
        (the local reduction kernel), then wait until the peers have finished pulling from this rank."""
        st = self._stream()
        e = self._epoch(CHAN_CE_IN)
        nbytes = shard_elems * 2
        _C.call("b2_signal_peers", self._pad_ptrs, self.rank, self.world, CHAN_CE_IN, e, st)
        _C.call("b2_stream_wait_pad", self._pad.data_ptr(), self.rank, self.world, CHAN_CE_IN, e, st)
        staging = self._ce_stage(nbytes * (self.world - 1))
        _C.call("b2_ce_pull_slices", staging.data_ptr(), fulls, nbytes, self.rank, self.world, st)
        _C.call("b2_signal_peers", self._pad_ptrs, self.rank, self.world, CHAN_CE_OUT, e, st)
        consume(staging)
        _C.call("b2_stream_wait_pad", self._pad.data_ptr(), self.rank, self.world, CHAN_CE_OUT, e, st)

    def _reduce_scatter(self, shard, full, average=True, max_blocks: int = 0, nvls: Optional[bool] = None):
        fulls = self._peer_array(full) if self.world > 1 else None
        if fulls is None or shard.dtype != torch.bfloat16:
            return TorchComm.reduce_scatter(self, shard, full, average)
        mc = self._mc_ptr(full) if (self.nvls if nvls is None else nvls) else 0
        if self._transport(nvls, bool(self._mc_ptr(full))) == "ce":
            n = shard.numel()
            local = full[self.rank * n:(self.rank + 1) * n]
            scale = (1.0 / self.world) if average else 1.0
            self._reduce_scatter_ce(fulls, n, local, lambda stg: _C.call(
                "b2_reduce_slices", shard.data_ptr(), local.data_ptr(), stg.data_ptr(), self.world - 1, n, scale,
                max_blocks, self._stream()))
            return
        if mc:
            _C.call("b2_nvls_reduce_scatter", shard.data_ptr(), mc, shard.numel(), self.rank, self.world,
                    (1.0 / self.world) if average else 1.0, self._pad_ptrs, CHAN_RS, self._epoch(CHAN_RS),
                    self._counter().data_ptr(), max_blocks or self.nvls_blocks, self._stream())
            return
        _C.call("b2_peer_reduce_scatter", shard.data_ptr(), fulls, shard.numel(), self.rank, self.world,
                (1.0 / self.world) if average else 1.0, self._pad_ptrs, CHAN_RS, self._epoch(CHAN_RS),
                self._counter().data_ptr(), max_blocks or self.rs_blocks, self._stream())

    def reduce_scatter_adamw(self, p, m, v, full_grad, lr, beta1, beta2, eps, wd, step, average=True,
                             max_blocks: int = 0, nvls: Optional[bool] = None) -> bool:
        """Fused gradient reduce-scatter + AdamW on the owned shard. Returns False when the buffers are not
        eligible (caller then runs the two steps separately)."""
        fulls = self._peer_array(full_grad) if self.world > 1 else None
        if fulls is None or p.dtype != torch.bfloat16:
            return False
        mc = self._mc_ptr(full_grad) if (self.nvls if nvls is None else nvls) else 0
        if self._transport(nvls, bool(self._mc_ptr(full_grad)), fused=True) == "ce":
            n = p.numel()
            local = full_grad[self.rank * n:(self.rank + 1) * n]
            scale = (1.0 / self.world) if average else 1.0
            self._reduce_scatter_ce(fulls, n, local, lambda stg: _C.call(
                "b2_reduce_adamw", p.data_ptr(), m.data_ptr(), v.data_ptr(), local.data_ptr(), stg.data_ptr(),
                self.world - 1, n, scale, lr, beta1, beta2, eps, wd, step, max_blocks or self.rsadam_blocks or 148,
                self._stream()))
            return True
        max_blocks = max_blocks or self.rsadam_blocks or (32 if mc else 148)
        if mc:
            _C.call("b2_nvls_reduce_scatter_adamw", p.data_ptr(), m.data_ptr(), v.data_ptr(), mc, p.numel(),
                    self.rank, self.world, (1.0 / self.world) if average else 1.0, lr, beta1, beta2, eps, wd,
                    step, self._pad_ptrs, CHAN_RS, self._epoch(CHAN_RS), self._counter().data_ptr(), max_blocks,
                    self._stream())
            return True
        _C.call("b2_peer_reduce_scatter_adamw", p.data_ptr(), m.data_ptr(), v.data_ptr(), fulls, p.numel(),
                self.rank, self.world, (1.0 / self.world) if average else 1.0, lr, beta1, beta2, eps, wd, step,
                self._pad_ptrs, CHAN_RS, self._epoch(CHAN_RS), self._counter().data_ptr(), max_blocks,
                self._stream())
        return True

    def _all_reduce(self, t, average=True, max_blocks: int = 64, nvls: Optional[bool] = None):
        bufs = self._peer_array(t) if self.world > 1 else None
        if bufs is None or t.dtype != torch.bfloat16 or t.numel() % (8 * self.world):
            return TorchComm.all_reduce(self, t, average)
        # the one-pass multimem all-reduce beats the two-phase peer kernel at every W measured
        mc = self._mc_ptr(t) if ((os.environ.get("B200TRAIN_NVLS", "auto") != "0") if nvls is None else nvls) else 0
        if mc:
            _C.call("b2_nvls_all_reduce", mc, t.numel(), self.rank, self.world,
                    (1.0 / self.world) if average else 1.0, self._pad_ptrs, CHAN_AR, self._epoch(CHAN_AR),
                    self._counter().data_ptr(), max_blocks, self._stream())
            return t
        _C.call("b2_peer_all_reduce", bufs, t.numel(), self.rank, self.world,
                (1.0 / self.world) if average else 1.0, self._pad_ptrs, CHAN_AR, self._epoch(CHAN_AR),
                self._counter().data_ptr(), max_blocks, self._stream())
        # the two-phase kernel returns while peers may still be pulling this rank's reduced slice: hold the
        # stream until every rank is out, so the caller may rewrite ``t`` as soon as the stream gets there
        _C.call("b2_peer_barrier", self._pad_ptrs, self.rank, self.world, CHAN_AR_EXIT,
                self._epoch(CHAN_AR_EXIT), self._stream())
        return t

    # ---- public surface: NVTX range + optional cross-check against NCCL --------------------------------
    def _check(self, what: str, got: torch.Tensor, want: torch.Tensor, tol: float):
        torch.cuda.synchronize(self.device)
        err = (got.float() - want.float()).abs()
        bound = tol * want.float().abs().clamp_min(1.0)
        bad = err > bound
        self.debug_checks += 1
        if bool(bad.any()):
            i = int(err.argmax())
            raise RuntimeError(f"peer collective mismatch in {what} on rank {self.rank}/{self.world}: element {i} "
                               f"is {float(got.flatten()[i])}, NCCL gives {float(want.flatten()[i])} "
                               f"({int(bad.sum())} of {got.numel()} elements outside tolerance)")

    def all_gather(self, full, shard, **kw):
        with nvtx.range("comm.all_gather"):
            self._all_gather(full, shard, **kw)
        if self.debug and self.world > 1:
            want = torch.empty_like(full)
            dist.all_gather_into_tensor(want, shard.clone(), group=self.group)
            self._check("all_gather", full, want, 0.0)

    def reduce_scatter(self, shard, full, average=True, **kw):
        snap = full.clone() if self.debug and self.world > 1 else None
        with nvtx.range("comm.reduce_scatter"):
            self._reduce_scatter(shard, full, average, **kw)
        if snap is not None:
            want = snap.float()                         # exact fp32 sum: NCCL's bf16 reduction rounds per hop
            dist.all_reduce(want, group=self.group)
            if average:
                want /= self.world
            n = shard.numel()
            self._check("reduce_scatter", shard, want[self.rank * n:(self.rank + 1) * n], 2 ** -6)

    def all_reduce(self, t, average=True, **kw):
        snap = t.clone() if self.debug and self.world > 1 else None
        with nvtx.range("comm.all_reduce"):
            out = self._all_reduce(t, average, **kw)
        if snap is not None:
            want = snap.float()
            dist.all_reduce(want, group=self.group)
            if average:
                want /= self.world
            self._check("all_reduce", t, want, 2 ** -6)
        return out
