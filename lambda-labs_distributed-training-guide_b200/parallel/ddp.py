"""Distributed data parallel with an optional ZeRO-1 (sharded optimizer state) update.

Reference: ``DistributedDataParallel(model, device_ids=[local_rank])`` + ``ZeroRedundancyOptimizer``
(``02-distributed-data-parallel/train_llm.py:65,81-83``):
  * parameters are broadcast from rank 0 at construction (N1),
  * gradients are averaged over ranks in buckets while backward is still running (N2). Here a bucket is a
    model unit's flat gradient buffer; its reduction is issued on the comm stream the moment the unit's
    backward finishes,
  * ``zero1=True``: each rank owns a contiguous 1/world slice of every unit: gradients are
    reduce-scattered, AdamW updates the owned slice, and the updated slices are all-gathered (the
    reference broadcasts per owner rank, N3). The sharded moments are saved per rank through the DCP
    path, which the reference skips because consolidation is slow (``02-.../README.md:352``),
  * ``no_sync()`` skips the reduction on non-boundary micro-steps
    (``related-topics/gradient-accumulation/README.md:28-38``).
"""
from __future__ import annotations

import contextlib
import os
from typing import Optional

import torch

from ..utils.flat import FlatUnit
from ..utils.schedule import make_lr_scheduler
from .base import Strategy
from .comm import Comm, make_comm
from .hooks import run_unit
from .optimizer import FlatAdamW


class DDP(Strategy):
    name = "ddp"

    def __init__(self, cfg, device, lr=3e-5, seed=0, dtype=torch.bfloat16, group=None,
                 comm: Optional[Comm] = None, zero1: bool = True, fuse_optimizer: bool = False,
                 prefer_peer: bool = True):
        super().__init__(cfg, device, lr, seed, dtype)
        self.comm = comm if comm is not None else make_comm(group, prefer_peer, self.device)
        self.W, self.R = self.comm.world, self.comm.rank
        self.zero1 = zero1 and self.W > 1
        self._zero1_mc = os.environ.get("B200TRAIN_ZERO1_MULTICAST", "1") != "0"
        self.fuse_optimizer = fuse_optimizer
        self.is_cuda = self.device.type == "cuda"
        self.sync_grads = True
        self.model = self._build_local_model()
        self.units = [FlatUnit(name, params, pad_to=8 * self.W,
                               alloc=lambda k: self.comm.empty(k, dtype, self.device))
                      for name, params in self.model.units()]
        for u in self.units:  # N1: every replica starts from rank 0's weights
            self.comm.broadcast(u.flat_param, 0)
        self.optimizer = FlatAdamW(self.units, lr=lr, shard=(self.R, self.W) if self.zero1 else None)
        self.lr_scheduler = make_lr_scheduler(self.optimizer, lr)
        self.comm_stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        self._param_events = {}   # unit -> event after its parameters were re-assembled / updated
        self._stepped = set()
        self.model.unit_runner = lambda idx, fn, *t: run_unit(self, idx, fn, *t)
        self.comm.barrier()

    @property
    def rank(self):
        return self.R

    @property
    def world_size(self):
        return self.W

    def tokens_per_step(self, batch_size, seq_len):
        return self.W * batch_size * seq_len

    def barrier(self):
        self.comm.barrier()

    # ---- helpers ------------------------------------------------------------------------------------
    def _record(self, stream=None):
        if not self.is_cuda:
            return None
        ev = torch.cuda.Event()
        ev.record(stream if stream is not None else torch.cuda.current_stream(self.device))
        return ev

    def _wait(self, ev, stream=None):
        if ev is not None and self.is_cuda:
            (stream if stream is not None else torch.cuda.current_stream(self.device)).wait_event(ev)

    @contextlib.contextmanager
    def _on_comm(self):
        if self.comm_stream is None:
            yield
        else:
            with torch.cuda.stream(self.comm_stream):
                yield

    # ---- unit callbacks -------------------------------------------------------------------------------
    def pre_forward(self, i):
        self._wait(self._param_events.pop(i, None))

    def post_forward(self, i):
        pass

    def pre_backward(self, i):
        self._wait(self._param_events.pop(i, None))

    def post_backward(self, i):
        u = self.units[i]
        u.finalize_grads()
        if not self.sync_grads:
            return
        ready = self._record()
        if self.W == 1:
            if self.fuse_optimizer:
                self._step_unit(i, ready)
            return
        with self._on_comm():
            self._wait(ready, self.comm_stream)
            if self.zero1:
                n = u.numel // self.W
                # the owned slice of the flat gradient receives the mean over ranks
                self.comm.reduce_scatter(u.flat_grad[self.R * n:(self.R + 1) * n], u.flat_grad, average=True)
            else:
                self.comm.all_reduce(u.flat_grad, average=True)
            done = self._record(self.comm_stream)
        if self.fuse_optimizer:
            self._step_unit(i, done)

    def _step_unit(self, i, after):
        if not self._stepped:
            self.optimizer.begin_step()
        self._stepped.add(i)
        with self._on_comm():
            self._wait(after, self.comm_stream)
            self.optimizer.step_unit(i, max_blocks=96 if self.is_cuda else 0, p_mc=self._param_mc(i))
            self._reassemble(i)
            self._param_events[i] = self._record(self.comm_stream)

    def _param_mc(self, i) -> int:
        """ZeRO-1 over symmetric memory with an NVSwitch: multicast address of this rank's slice of unit ``i``'s
        flat parameters (0: none -- the slices are gathered after the update instead)."""
        if not (self.zero1 and self.is_cuda and self._zero1_mc):
            return 0
        mc = self.comm._mc_ptr(self.units[i].flat_param) if hasattr(self.comm, "_mc_ptr") else 0
        if not mc:
            return 0
        u = self.units[i]
        return mc + self.R * (u.numel // self.W) * u.flat_param.element_size()

    def _reassemble(self, i):
        """ZeRO-1: make every replica hold the updated parameter slices (N3). With a multicast mapping the
        AdamW kernel has already stored its slice into every rank's copy (``multimem.st``) and a device-side
        barrier is all that is left; otherwise the slices are gathered by the copy engines."""
        if not self.zero1:
            return
        u = self.units[i]
        self.comm.stream_barrier()  # every owner finished updating its slice
        if not self._param_mc(i):
            self.comm.all_gather_inplace(u.flat_param)

    # ---- phases ---------------------------------------------------------------------------------------
    def backward(self, loss):
        loss.backward()
        self.post_backward(0)  # the embedding unit has no differentiable input to fire it

    @contextlib.contextmanager
    def no_sync(self):
        old = self.sync_grads
        self.sync_grads = False
        try:
            yield
        finally:
            self.sync_grads = old

    def optimizer_step(self):
        if self.is_cuda:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        if not self._stepped:
            self.optimizer.begin_step()
        for i in range(len(self.units)):
            if i not in self._stepped:
                self.optimizer.step_unit(i, p_mc=self._param_mc(i))
                self._reassemble(i)
        self._stepped.clear()
        self._param_events.clear()
        self.lr_scheduler.step()
        if self.is_cuda:
            self.comm_stream.wait_stream(torch.cuda.current_stream(self.device))

    # ---- checkpoint surface (moments are sharded under zero1) -------------------------------------------
    def _unit_full_flat(self, i, kind):
        if kind == "param" or not self.zero1:
            return super()._unit_full_flat(i, kind)
        shard = (self.optimizer.exp_avg if kind == "exp_avg" else self.optimizer.exp_avg_sq)[i]
        full = torch.empty(self.units[i].numel, dtype=self.dtype, device=self.device)
        self.comm.all_gather(full, shard)
        return full

    def _unit_store_flat(self, i, kind, full):
        if kind == "param" or not self.zero1:
            return super()._unit_store_flat(i, kind, full)
        shard = (self.optimizer.exp_avg if kind == "exp_avg" else self.optimizer.exp_avg_sq)[i]
        n = shard.numel()
        shard.copy_(full[self.R * n:(self.R + 1) * n])
