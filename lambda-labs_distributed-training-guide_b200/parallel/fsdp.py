"""Fully sharded data parallel (ZeRO-3) over flat per-unit shards.

Capabilities of the reference's FSDP chapters:
  * FULL_SHARD of parameters, gradients and optimizer state, one flat shard per unit
    (``04-fully-sharded-data-parallel/train_llm.py:84-95``); units = embedding, each decoder layer,
    final norm + head (the transformer wrap policy of ``05-training-llama-405b/train_llm.py:101-107``;
    ``--numel-to-wrap`` merges small units the way ``size_based_auto_wrap_policy`` does),
  * construction without materialising the full model: every unit is initialised into a transient
    buffer and only the local shard is kept (meta-device init + ``sync_module_states``:
    ``04-.../train_llm.py:76-79,90``),
  * all-gather before forward / backward of a unit with the next unit prefetched on a comm stream,
    reduce-scatter (mean) of the unit's gradients right after its backward,
  * ``cpu_offload``: parameter shards, gradient shards and optimizer state live in pinned host memory;
    H2D/D2H copies run on a side stream and AdamW runs on the CPU shard
    (``CPUOffload(offload_params=True)``: ``04-.../train_llm.py:94``, ``05-.../README.md:147-169``),
  * activation checkpointing per decoder layer (``05-.../train_llm.py:124-126``),
  * ``no_sync`` gradient accumulation (``related-topics/gradient-accumulation/README.md:28-38``).

With ``fuse_optimizer=True`` the sharded AdamW update of a unit is launched as soon as its reduced
gradient shard exists, i.e. inside backward, overlapping the remaining backward compute.
"""
from __future__ import annotations

import contextlib
from typing import List, Optional

import torch

from ..models import build_model
from ..utils.flat import FlatUnit
from ..utils.schedule import make_lr_scheduler
from .base import Strategy
from .comm import Comm, make_comm
from .hooks import run_unit
from .optimizer import FlatAdamW


import os

# blocks given to an AdamW launched beside backward compute (co-resident with the persistent GEMM CTAs)
_FUSED_ADAMW_BLOCKS = int(os.environ.get("B200TRAIN_FUSED_ADAMW_BLOCKS", "296"))


class _BufferPool:
    """A few full-size buffers with a fixed unit -> slot mapping (unit ``i`` always lands at the same
    address, so tensors autograd saved in forward stay valid when the unit is re-gathered in backward).
    Each slot remembers the CUDA event after which it may be overwritten."""

    def __init__(self, n: int, numel: int, dtype, device, dedicated_first: bool = False, alloc=None):
        alloc = alloc or (lambda k: torch.empty(k, dtype=dtype, device=device))
        self.bufs = [alloc(numel) for _ in range(n)]
        self.events = [None] * n
        self.owner = [None] * n
        self.n = n
        self.dedicated_first = dedicated_first  # slot 0 is reserved for unit 0 (tied embedding)

    def slot_of(self, unit: int) -> int:
        if self.dedicated_first:
            return 0 if unit == 0 else 1 + (unit - 1) % (self.n - 1)
        return unit % self.n

    def acquire(self, unit: int):
        i = self.slot_of(unit)
        if self.owner[i] is not None:
            raise RuntimeError(f"FSDP buffer slot {i} still owned by unit {self.owner[i]} (wanted by {unit})")
        self.owner[i] = unit
        return i, self.bufs[i], self.events[i]

    def release(self, i: int, event=None):
        self.events[i] = event
        self.owner[i] = None


class _Gathered:
    """One in-flight / resident all-gather of a unit: pool slot, completion event, the full view and, for
    the flag-gated (fused) gather, the event after the own-shard copy plus the arrival flags and epoch."""

    __slots__ = ("slot", "ev", "full", "ev_local", "flags", "epoch")

    def __init__(self, slot, ev, full, ev_local=None, flags=None, epoch=0):
        self.slot, self.ev, self.full, self.ev_local, self.flags, self.epoch = slot, ev, full, ev_local, flags, epoch


class FSDP(Strategy):
    name = "fsdp"

    def __init__(self, cfg, device, lr=3e-5, seed=0, dtype=torch.bfloat16, group=None,
                 comm: Optional[Comm] = None, cpu_offload: bool = False, activation_checkpointing: bool = False,
                 reshard_after_forward: bool = True, fuse_optimizer: bool = False, numel_to_wrap: int = 0,
                 prefer_peer: bool = True, sync_module_states: bool = False):
        super().__init__(cfg, device, lr, seed, dtype)
        # N1 (``sync_module_states=True`` of the reference, 04-.../train_llm.py:87-95): every unit is broadcast
        # from the first rank of the group while it is being materialised. Off by default: all ranks initialise
        # from the same seed, which gives the same parameters without the traffic; switch it on when the ranks
        # were (or may have been) seeded differently.
        self.sync_module_states = sync_module_states
        self.comm = comm if comm is not None else make_comm(group, prefer_peer, self.device)
        self.W, self.R = self.comm.world, self.comm.rank
        self.cpu_offload = cpu_offload
        self.act_ckpt = activation_checkpointing
        # fuse_optimizer: False | True / "kernel" (gradient reduce-scatter + AdamW on the owned shard as one
        # fused operation when the communicator provides it: copy-engine pull of the peers' slices, one local
        # kernel that reduces and updates) | "stream" (AdamW kernel queued right behind the reduce-scatter) |
        # "auto": "kernel" up to 4 ranks, "stream" from 8 (measured per Llama-8B step, kernel / stream:
        # 2 GPUs 197.8-199.6 / 202.7 ms, 4 GPUs 198.4 / 198.6 ms, 8 GPUs 201.0 / 196.9 ms: with 7 staged
        # slices per unit the in-switch reduction that feeds a separate AdamW moves fewer HBM bytes)
        if fuse_optimizer == "auto":
            fuse_optimizer = "stream" if self.W >= 8 else "kernel"
        self.fuse_optimizer = fuse_optimizer and not cpu_offload
        self.fuse_in_kernel = fuse_optimizer in (True, "kernel")
        self.is_cuda = self.device.type == "cuda"
        self.numel_to_wrap = numel_to_wrap
        self.sync_grads = True
        # resident mode: one process, nothing offloaded -> parameters and gradients stay in their flat
        # buffers and the gather / scatter machinery is bypassed
        self.resident = self.W == 1 and not cpu_offload
        self.reshard = reshard_after_forward and not self.resident
        tied = cfg.tie_word_embeddings

        # ---- model skeleton on the meta device; units define the flat layout -------------------
        self.model = self._build_meta_model(cfg, dtype)
        if numel_to_wrap and hasattr(self.model, "group_layers"):
            self.model.group_layers(numel_to_wrap)
        self.units: List[_Layout] = [_Layout(name, params, self.W) for name, params in self.model.units()]
        n_units = len(self.units)
        max_numel = max(u.numel for u in self.units)
        if self.resident:
            self.param_pool = self.grad_pool = None
        else:
            # transient full buffers: current + prefetched (+ the tied embedding kept resident);
            # without resharding every unit keeps its gathered copy for the whole step
            # offload: gathers start with a PCIe copy of the shard, so they are prefetched two units ahead
            self.prefetch = 2 if cpu_offload else 1
            self.param_pool = _BufferPool((2 + self.prefetch + int(tied)) if self.reshard else n_units, max_numel,
                                          dtype, self.device, dedicated_first=tied and self.reshard)
            # gradient buffers are read by the peers' reduce-scatter: allocate them through the comm
            self.grad_pool = _BufferPool(2 + int(tied), max_numel, dtype, self.device, dedicated_first=tied,
                                         alloc=(None if cpu_offload else
                                                lambda k: self.comm.empty(k, dtype, self.device)))
        host = dict(device="cpu", pin_memory=self.is_cuda) if cpu_offload else dict(device=self.device)
        self.shard_param: List[torch.Tensor] = []
        self.shard_grad: List[torch.Tensor] = []
        swap = {}
        for ui, u in enumerate(self.units):
            n = u.numel // self.W
            full = (torch.empty(u.numel, dtype=dtype, device=self.device) if self.resident
                    else self.param_pool.bufs[0][:u.numel])
            swap.update(u.materialize(self.model, full, self._init_seed(seed), ui, cfg))
            if self.sync_module_states:
                self._sync_unit(full)
            if self.resident:
                self.shard_param.append(full)
                self.shard_grad.append(torch.zeros(u.numel, dtype=dtype, device=self.device))
            else:
                shard = torch.empty(n, dtype=dtype, **host) if cpu_offload else self.comm.empty(n, dtype, self.device)
                shard.copy_(full[self.R * n:(self.R + 1) * n])
                self.shard_param.append(shard)
                self.shard_grad.append(torch.zeros(n, dtype=dtype, **host))
        for mod in self.model.modules():
            for pname, old in list(mod._parameters.items()):
                if old is not None and id(old) in swap:
                    mod._parameters[pname] = swap[id(old)]
        if self.resident:
            for u, g in zip(self.units, self.shard_grad):
                u.assign_grads(g)
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        self._gathered = {}       # unit -> _Gathered
        # all-gather fused into the consuming GEMMs: the gather is issued in chunks with an arrival flag each
        # and the linears of the unit wait for their own weight rows inside the GEMM (ops/linear.py,
        # b2_gemm_bf16_bwait) instead of the compute stream waiting for the whole unit
        self.fused_ag = (self.supports_fused_ag and self.is_cuda and self.W > 1 and not cpu_offload
                         and not self.resident and type(self.comm).__name__ == "PeerComm" and dtype == torch.bfloat16
                         and os.environ.get("B200TRAIN_FSDP_FUSED_AG", "1") != "0")
        if self.fused_ag:
            self._ag_flags = torch.zeros((self.param_pool.n, 64), dtype=torch.int32, device=self.device)
            self._ag_epochs = [0] * self.param_pool.n
            sub = max(1, min(16, 64) // self.W)
            for u in self.units:
                u.sub = sub
                while u.sub > 1 and (u.numel // self.W) % (u.sub * 8):
                    u.sub //= 2
        self._grad_slot = {}      # unit -> pool slot of its full gradient buffer
        self._opt_events = {}     # unit -> event after a side-stream optimizer update of its shard
        self._local_acc = {}      # unit -> unreduced gradient accumulator (no_sync)
        self._stepped = set()     # units already updated inside this backward (fuse_optimizer)
        self._began = False       # optimizer.begin_step() already called for the step in flight
        # two side streams: parameter gathers (copy engines) must not queue behind a unit's reduce-scatter and
        # optimizer update, or the prefetch of the next unit starts late
        self.comm_stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        self.ag_stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        # CPU offload: the host AdamW of a unit starts on a worker thread as soon as its gradient shard
        # has landed in pinned memory, overlapping the rest of backward (the reference steps after it)
        self._cpu_worker = None
        self._cpu_jobs = []
        if cpu_offload and self.is_cuda:
            from concurrent.futures import ThreadPoolExecutor

            # two workers: one can sit in the event wait of the next unit's gradient copy while the other updates
            self._cpu_worker = ThreadPoolExecutor(max_workers=2)
            # torchrun pins every rank to OMP_NUM_THREADS=1: give the host AdamW its share of the cores instead
            local = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
            want = int(os.environ.get("B200TRAIN_CPU_ADAMW_THREADS", "0")) or max(1, (os.cpu_count() or 8) // local - 2)
            try:
                from .. import _C

                lib = _C.load()
                if hasattr(lib, "b2_cpu_set_threads"):
                    lib.b2_cpu_set_threads(want)
            except Exception:
                pass
        self.model.unit_runner = self._runner
        self.model.train()

        self.optimizer = FlatAdamW(self.units, lr=lr, shard=(self.R, self.W),
                                   moments_device=("cpu" if cpu_offload else self.device),
                                   pin=self.is_cuda and cpu_offload)
        self.lr_scheduler = make_lr_scheduler(self.optimizer, lr)
        self.comm.barrier()

    # ---- extension points (the 2D strategy plugs a tensor-parallel model in here) ----------------------
    def _build_meta_model(self, cfg, dtype):
        return build_model(cfg, dtype=dtype, device="meta", init=False)

    def _init_seed(self, seed):
        return seed

    def _sync_unit(self, full):
        """Make every rank of the sharding group hold rank 0's freshly initialised unit."""
        self.comm.broadcast(full, 0)

    def pre_reduce(self, i):
        """Called when unit ``i``'s gradients are complete, before they are reduced over this group."""

    # False for strategies whose linears are not ops.linear (the tensor-parallel fused linears of 2D)
    supports_fused_ag = True

    # set by strategies that reduce the gradient shard over a second group (hybrid shard): the fused
    # reduce-scatter + AdamW kernel is then bypassed, the update must see the fully reduced shard
    has_post_reduce = False

    def post_reduce(self, i, shard):
        """Called on the comm stream with unit ``i``'s reduced gradient shard, before it is consumed."""

    # ------------------------------------------------------------------------------------------------
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

    # ---- stream / event helpers -----------------------------------------------------------------------
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
    def _on(self, stream):
        if stream is None:
            yield
        else:
            with torch.cuda.stream(stream):
                yield

    # ---- gather / release ---------------------------------------------------------------------------
    def _gather(self, i: int):
        """Issue the all-gather of unit ``i`` on the comm stream (no-op if it is already in flight)."""
        if self.resident or i in self._gathered or i < 0 or i >= len(self.units):
            return
        u = self.units[i]
        slot, buf, free_ev = self.param_pool.acquire(i)
        full = buf[:u.numel]
        n = u.numel // self.W
        if self.fused_ag and u.prefix < u.numel:
            self._ag_epochs[slot] += 1
            epoch, flags = self._ag_epochs[slot], self._ag_flags[slot]
            with self._on(self.ag_stream):
                self._wait(free_ev, self.ag_stream)
                # own shard first (a copy KERNEL: consumers order themselves behind it with an event), then the
                # peers' chunks by copy engine, a flag written by the stream after each
                self.comm.all_gather_flagged(full, self.shard_param[i], flags, epoch, u.sub, part=0)
                ev_local = self._record(self.ag_stream)
                self.comm.all_gather_flagged(full, self.shard_param[i], flags, epoch, u.sub, part=1)
                ev = self._record(self.ag_stream)
            self._gathered[i] = _Gathered(slot, ev, full, ev_local, flags, epoch)
            return
        with self._on(self.ag_stream):
            self._wait(free_ev, self.ag_stream)
            if self.cpu_offload:
                src = full[self.R * n:(self.R + 1) * n]
                src.copy_(self.shard_param[i], non_blocking=True)   # pinned H2D on the side stream
            else:
                src = self.shard_param[i]
            if self.W > 1:
                self.comm.all_gather(full, src)
            ev = self._record(self.ag_stream)
        self._gathered[i] = _Gathered(slot, ev, full)

    def _ensure(self, i: int):
        if self.resident:
            self._wait(self._opt_events.pop(i, None))
            return
        self._gather(i)
        g = self._gathered[i]
        u = self.units[i]
        if g.flags is None:
            self._wait(g.ev)
            u.repoint(g.full)
            return
        # flag-gated gather: behind the own-shard copy, and behind the chunks that hold the parameters no
        # GEMM consumes (norm weights, biases: laid out first in the unit); the linears wait in-kernel
        from ..ops.linear import GatherWait

        self._wait(g.ev_local)
        n = u.numel // self.W
        chunk = n // u.sub
        if u.prefix > 0:
            _C_call_wait(g.flags, 0, (u.prefix + chunk - 1) // chunk, g.epoch, self.device)
        u.repoint(g.full)
        fp = g.flags.data_ptr()
        for p, o in zip(u.params, u.offsets):
            p._b2_ag = (GatherWait(fp, g.epoch, o, chunk, self.W * u.sub, u.sub, self.R, self.W)
                        if getattr(p, "_b2_linear_only", False) else None)

    def _release(self, i: int):
        if self.resident or i not in self._gathered:
            return
        g = self._gathered.pop(i)
        self.param_pool.release(g.slot, self._record())

    def _tied_head(self, i: int) -> bool:
        return self.cfg.tie_word_embeddings and i == len(self.units) - 1

    # ---- unit callbacks (see hooks.run_unit) ----------------------------------------------------------
    def _runner(self, idx, fn, *tensors):
        ckpt = self.act_ckpt and 1 <= idx <= self.cfg.num_hidden_layers
        return run_unit(self, idx, fn, *tensors, activation_checkpoint=ckpt)

    def pre_forward(self, i):
        self._ensure(i)
        if self._tied_head(i):
            self._ensure(0)
        for k in range(1, getattr(self, "prefetch", 1) + 1):
            self._gather(i + k)

    def post_forward(self, i):
        last = i == len(self.units) - 1
        keep = (self.cfg.tie_word_embeddings and i == 0) or (last and torch.is_grad_enabled())
        if self.reshard and not keep:
            self._release(i)

    def _assign_grad_buffer(self, i):
        if self.resident or i in self._grad_slot:
            return
        u = self.units[i]
        slot, buf, free_ev = self.grad_pool.acquire(i)
        self._wait(free_ev)
        u.assign_grads(buf[:u.numel])
        self._grad_slot[i] = slot

    def pre_backward(self, i):
        self._ensure(i)
        self._assign_grad_buffer(i)
        if self._tied_head(i):
            self._ensure(0)
            self._assign_grad_buffer(0)
        for k in range(1, getattr(self, "prefetch", 1) + 1):
            self._gather(i - k)

    def post_backward(self, i):
        u = self.units[i]
        u.finalize_grads()
        self.pre_reduce(i)
        if self.resident:
            if self.fuse_optimizer and self.sync_grads:
                self._fused_step(i, self._record())
            return
        if self.sync_grads:
            self._reduce(i)
        else:
            self._accumulate_local(i)
        self._release(i)

    # ---- gradient reduction ---------------------------------------------------------------------------
    def _accumulate_local(self, i):
        """no_sync micro-step: park the unreduced gradient of the unit until the synchronising step."""
        u = self.units[i]
        if i in self._local_acc:
            self._local_acc[i].add_(u.flat_grad)
        else:
            self._local_acc[i] = u.flat_grad.clone()
        self.grad_pool.release(self._grad_slot.pop(i), self._record())

    def _reduce(self, i):
        u = self.units[i]
        n = u.numel // self.W
        full = u.flat_grad
        if i in self._local_acc:
            full.add_(self._local_acc.pop(i))
        ready = self._record()
        slot = self._grad_slot.pop(i)
        if self.fuse_optimizer and self.fuse_in_kernel and self.W > 1 and not self.has_post_reduce:
            # one kernel: pull + reduce the peers' gradient slices and apply AdamW to the owned shard
            with self._on(self.comm_stream):
                self._wait(ready, self.comm_stream)
                self._begin_once()
                g = self.optimizer.param_groups[0]
                if self.comm.reduce_scatter_adamw(self.shard_param[i], self.optimizer.exp_avg[i],
                                                  self.optimizer.exp_avg_sq[i], full, g["lr"], g["betas"][0],
                                                  g["betas"][1], g["eps"], g["weight_decay"], self.optimizer._step):
                    self._stepped.add(i)
                    self.grad_pool.release(slot, self._record(self.comm_stream))
                    return
        with self._on(self.comm_stream):
            self._wait(ready, self.comm_stream)
            if self.cpu_offload:
                if self.W > 1:
                    dev = torch.empty(n, dtype=full.dtype, device=self.device)
                    self.comm.reduce_scatter(dev, full, average=True)
                else:
                    dev = full[:n]
                self.post_reduce(i, dev)
                self.shard_grad[i].copy_(dev, non_blocking=True)      # D2H into pinned memory
            else:
                self.comm.reduce_scatter(self.shard_grad[i], full, average=True)
                self.post_reduce(i, self.shard_grad[i])
            done = self._record(self.comm_stream)
        if self._cpu_worker is not None:
            self._submit_cpu_step(i, done)
        self.grad_pool.release(slot, done)
        if self.fuse_optimizer:
            self._fused_step(i, done)

    def _submit_cpu_step(self, i, landed):
        self._begin_once()
        self._stepped.add(i)

        def job():
            landed.synchronize()          # the D2H copy of this unit's gradient shard has finished
            self.optimizer.step_unit(i, grad=self.shard_grad[i], param=self.shard_param[i])

        self._cpu_jobs.append(self._cpu_worker.submit(job))

    def _fused_step(self, i, after_event):
        """Sharded AdamW for unit ``i`` on the side stream as soon as its gradient shard is final."""
        self._begin_once()
        self._stepped.add(i)
        with self._on(self.comm_stream):
            self._wait(after_event, self.comm_stream)
            self.optimizer.step_unit(i, grad=self.shard_grad[i], param=self.shard_param[i],
                                     max_blocks=(_FUSED_ADAMW_BLOCKS if self.is_cuda else 0))
            if self.resident:
                self._opt_events[i] = self._record(self.comm_stream)

    def _begin_once(self):
        """Advance the optimizer's step counter exactly once per optimizer step, whichever path (fused kernel,
        side-stream AdamW, host AdamW, end-of-step sweep) reaches a unit first."""
        if not self._began:
            self.optimizer.begin_step()
            self._began = True

    # ---- phases ---------------------------------------------------------------------------------------
    def forward(self, batch):
        if not self.resident and self.W > 1:
            # every rank's parameter shards are final (optimizer of the previous step) before any pull
            with self._on(self.comm_stream):
                self.comm.stream_barrier()
                self._wait(self._record(self.comm_stream), self.ag_stream)
        return self.model(**batch)

    def backward(self, loss):
        loss.backward()
        self.post_backward(0)  # the embedding unit has no differentiable input to fire it
        for i in list(self._gathered):
            self._release(i)

    def zero_grad(self):
        for u in self.units:
            u.zero_grad()

    @contextlib.contextmanager
    def no_sync(self):
        """Skip the gradient reduction for micro-steps inside the context (gradient accumulation,
        ``related-topics/gradient-accumulation/README.md:28-38``)."""
        old = self.sync_grads
        self.sync_grads = False
        try:
            yield
        finally:
            self.sync_grads = old

    def optimizer_step(self):
        if self.is_cuda:
            if self.cpu_offload:
                for j in self._cpu_jobs:          # host updates started during backward
                    j.result()
                self._cpu_jobs.clear()
                self.comm_stream.synchronize()   # gradient shards must have landed in host memory
            else:
                torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        self._begin_once()
        for i in range(len(self.units)):
            if i not in self._stepped:
                self.optimizer.step_unit(i, grad=self.shard_grad[i], param=self.shard_param[i])
        self._stepped.clear()
        self._began = False
        self.lr_scheduler.step()
        if self.is_cuda:
            self.comm_stream.wait_stream(torch.cuda.current_stream(self.device))

    # ---- state ----------------------------------------------------------------------------------------
    @torch.no_grad()
    def gather_full_state_dict(self):
        """HF-named full (unsharded) parameters on every rank -- for tests and export."""
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        keep = []
        for i, u in enumerate(self.units):
            if self.resident:
                continue
            full = torch.empty(u.numel, dtype=self.dtype, device=self.device)
            n = u.numel // self.W
            full[self.R * n:(self.R + 1) * n].copy_(self.shard_param[i])
            if self.W > 1:
                self.comm.all_gather(full, full[self.R * n:(self.R + 1) * n].clone())
            u.repoint(full)
            keep.append(full)
        return {k: v.clone() for k, v in self.model.hf_state_dict().items()}

    def model_state_dict(self):
        return self.gather_full_state_dict()

    # ---- checkpoint surface --------------------------------------------------------------------------
    def _shard_of(self, i, kind):
        if kind == "param":
            return self.shard_param[i]
        return (self.optimizer.exp_avg if kind == "exp_avg" else self.optimizer.exp_avg_sq)[i]

    def _unit_full_flat(self, i, kind):
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        shard = self._shard_of(i, kind)
        if self.resident:
            return shard
        u = self.units[i]
        n = u.numel // self.W
        full = torch.empty(u.numel, dtype=self.dtype, device=self.device)
        mine = shard.to(self.device)
        if self.W > 1:
            self.comm.all_gather(full, mine)
        else:
            full.copy_(mine)
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        return full

    def _unit_store_flat(self, i, kind, full):
        shard = self._shard_of(i, kind)
        if self.resident:
            shard.copy_(full)
            return
        n = self.units[i].numel // self.W
        shard.copy_(full[self.R * n:(self.R + 1) * n])


    # ---- checkpoint surface without full-size gathers ---------------------------------------------------
    # Reference: get_state_dict(..., StateDictOptions(full_state_dict=False)) + DCP save / load of the sharded
    # containers (04-fully-sharded-data-parallel/train_llm.py:122,132-154,249-263), where FSDP1's flat-parameter
    # runtime re-chunks into ShardedTensors internally.
    # torch.distributed.checkpoint wants, from rank r, the r-th dim-0 chunk of every named tensor; the strategy
    # owns the r-th slice of every flat unit. The two partitions cover the same bytes, so re-chunking is an
    # exchange in which every element moves at most once: rank k sends to rank r the overlap of flat slice k with
    # the chunks of rank r (one all-to-all per unit and kind, numel/W elements per rank), instead of gathering the
    # whole unit on every rank and discarding (W-1)/W of it.
    def _chunk_plan(self, i):
        """``[(name, shape, flat offset)]`` of unit ``i``'s named tensors in flat order, or None when a named tensor
        is not one contiguous range of the flat unit (then the gather path is used)."""
        plans = self.__dict__.setdefault("_chunk_plans", {})
        if i not in plans:
            probe = self._names_probe(i)
            base, es = probe.data_ptr(), probe.element_size()
            plan = []
            for name, v in self._named_views(i, probe).items():
                if not v.is_contiguous():
                    plan = None
                    break
                plan.append((name, tuple(v.shape) if v.dim() else (1,), (v.data_ptr() - base) // es))
            plans[i] = sorted(plan, key=lambda t: t[2]) if plan is not None else None
        return plans[i]

    def _xchg(self):
        """Ranks taking part in the re-chunking exchange: ``(group, n, me, slice_of, same_replica)``. Rank r of n
        holds DCP chunk r of every tensor and flat slice ``slice_of(r)``; ``same_replica(a, b)`` says whether a's
        slice may serve b's chunk (hybrid shard: every replica holds every slice, only b's own replica sends)."""
        return getattr(self.comm, "group", None), self.W, self.R, (lambda r: r), (lambda a, b: True)

    def _chunk_range(self, shape, off, r, n):
        """Flat range of the dim-0 chunk that rank ``r`` of ``n`` holds (torch's ``Shard(0)`` split: ceil)."""
        rows = shape[0]
        rowsz = 1
        for d in shape[1:]:
            rowsz *= d
        per = (rows + n - 1) // n
        lo = min(r * per, rows)
        hi = min(lo + per, rows)
        return off + lo * rowsz, off + hi * rowsz, (hi - lo,) + tuple(shape[1:])

    def _overlaps(self, i, chunk_rank, n, slice_index):
        """``[(j, lo, hi)]``: flat ranges lying both in flat slice ``slice_index`` and in a chunk of ``chunk_rank``."""
        per = self.units[i].numel // self.W
        s_lo, s_hi = slice_index * per, (slice_index + 1) * per
        out = []
        for j, (_, shape, off) in enumerate(self._chunk_plan(i)):
            a, b, _ = self._chunk_range(shape, off, chunk_rank, n)
            lo, hi = max(a, s_lo), min(b, s_hi)
            if lo < hi:
                out.append((j, lo, hi))
        return out

    def _supports_chunk_exchange(self, i):
        return not self.resident and self._xchg()[1] > 1 and self._chunk_plan(i) is not None

    def _all_to_all(self, group, out, src, outs, ins):
        import torch.distributed as dist

        dist.all_to_all_single(out, src, list(outs), list(ins), group=group)

    @torch.no_grad()
    def unit_named_chunks(self, i: int, kind: str):
        """``{name: (local dim-0 chunk, global shape)}`` of unit ``i`` (``kind``: param / exp_avg / exp_avg_sq) for this
        rank, or None when the exchange does not apply (single rank, non-contiguous named tensors)."""
        if not self._supports_chunk_exchange(i):
            return None
        if self.is_cuda:
            torch.cuda.synchronize(self.device)
        group, n, me, slice_of, same_replica = self._xchg()
        plan = self._chunk_plan(i)
        per = self.units[i].numel // self.W
        my_lo = slice_of(me) * per
        mine = self._shard_of(i, kind).to(self.device)
        send, ins = [], []
        for r in range(n):
            ov = self._overlaps(i, r, n, slice_of(me)) if same_replica(me, r) else []
            parts = [mine[lo - my_lo:hi - my_lo] for _, lo, hi in ov]
            ins.append(sum(p.numel() for p in parts))
            send += parts
        recv_plan = [self._overlaps(i, me, n, slice_of(k)) if same_replica(k, me) else [] for k in range(n)]
        outs = [sum(hi - lo for _, lo, hi in ov) for ov in recv_plan]
        got = torch.empty(sum(outs), dtype=mine.dtype, device=self.device)
        src = torch.cat(send) if send else mine[:0]
        self._all_to_all(group, got, src, outs, ins)
        chunks, flat_lo = {}, {}
        for j, (name, shape, off) in enumerate(plan):
            a, _, cshape = self._chunk_range(shape, off, me, n)
            chunks[name] = (torch.empty(cshape, dtype=mine.dtype, device=self.device), torch.Size(shape))
            flat_lo[j] = a
        pos = 0
        for ov in recv_plan:
            for j, lo, hi in ov:
                chunks[plan[j][0]][0].view(-1)[lo - flat_lo[j]:hi - flat_lo[j]].copy_(got[pos:pos + hi - lo])
                pos += hi - lo
        return chunks

    @torch.no_grad()
    def unit_store_chunks(self, i: int, kind: str, chunks) -> bool:
        """Inverse of :meth:`unit_named_chunks`: ``chunks[name]`` is this rank's dim-0 chunk of every named tensor
        of unit ``i``; every element travels to the rank (under hybrid shard: the ranks) whose flat slice holds it."""
        if not self._supports_chunk_exchange(i):
            return False
        group, n, me, slice_of, _ = self._xchg()
        plan = self._chunk_plan(i)
        per = self.units[i].numel // self.W
        my_lo = slice_of(me) * per
        flat_lo = {j: self._chunk_range(shape, off, me, n)[0] for j, (_, shape, off) in enumerate(plan)}
        local = {j: chunks[name].to(device=self.device, dtype=self.dtype).reshape(-1)
                 for j, (name, _, _) in enumerate(plan)}
        send, ins = [], []
        for k in range(n):
            parts = [local[j][lo - flat_lo[j]:hi - flat_lo[j]] for j, lo, hi in self._overlaps(i, me, n, slice_of(k))]
            ins.append(sum(p.numel() for p in parts))
            send += parts
        recv_plan = [self._overlaps(i, r, n, slice_of(me)) for r in range(n)]
        outs = [sum(hi - lo for _, lo, hi in ov) for ov in recv_plan]
        got = torch.empty(sum(outs), dtype=self.dtype, device=self.device)
        src = torch.cat(send) if send else got[:0]
        self._all_to_all(group, got, src, outs, ins)
        new = torch.zeros(per, dtype=self.dtype, device=self.device)   # gaps between tensors (alignment) stay zero
        pos = 0
        for ov in recv_plan:
            for _, lo, hi in ov:
                new[lo - my_lo:hi - my_lo].copy_(got[pos:pos + hi - lo])
                pos += hi - lo
        self._shard_of(i, kind).copy_(new)
        return True


def _C_call_wait(flags, first, count, epoch, device):
    from .. import _C

    _C.call("b2_stream_wait_flags", flags.data_ptr(), first, count, epoch,
            torch.cuda.current_stream(device).cuda_stream)


class _Layout(FlatUnit):
    """A FlatUnit whose storage is transient: ``repoint``/``assign_grads`` attach it to pool buffers."""

    def __init__(self, name, params, world):
        # layout only (parameters are on the meta device at this point)
        self.name = name
        self.params = []
        seen = set()
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        from ..utils.flat import ALIGN, _round_up

        # parameters that only GEMMs read (tagged by the model) go last: whatever precedes them -- norm
        # weights, biases, embeddings -- is what the compute stream has to wait for before the unit runs,
        # the linears wait for their own rows inside the GEMM (fused all-gather)
        self.params.sort(key=lambda p: bool(getattr(p, "_b2_linear_only", False)))
        self.offsets = []
        off = 0
        self.prefix = None
        for p in self.params:
            if self.prefix is None and getattr(p, "_b2_linear_only", False):
                self.prefix = off
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.numel = _round_up(off, ALIGN * world)
        if self.prefix is None:
            self.prefix = self.numel
        self.sub = 1
        self.dtype = self.params[0].dtype
        self.device = None
        self.flat_param = None
        self.flat_grad = None
        self.direct = False

    def materialize(self, model, full, seed, unit_index, cfg):
        """Initialise this unit's parameters into ``full`` (HF ``_init_weights`` rules), identically on
        every rank (same generator seed), which stands in for rank-0 init + broadcast."""
        self.device = full.device
        self.direct = full.device.type == "cuda" and full.dtype == torch.bfloat16
        full.zero_()
        gen = torch.Generator(device=full.device).manual_seed(seed * 100003 + unit_index)
        names = {id(p): n for n, p in model.named_parameters()}
        for p, o in zip(self.params, self.offsets):
            view = full[o:o + p.numel()].view(p.shape)
            name = names.get(id(p), "")
            if name.endswith("bias"):
                view.zero_()
            elif "norm" in name or ".ln_" in name or "ln_f" in name:
                view.fill_(1.0)
            else:
                view.normal_(0.0, cfg.initializer_range, generator=gen)
        self.flat_param = full
        # meta parameters cannot adopt real storage: create real Parameters (views of the flat buffer)
        # and let the caller swap them into the module tree (ties are preserved by identity)
        swap = {}
        new_params = []
        for p, o in zip(self.params, self.offsets):
            # (.data assignment: the Parameter must not share the pool buffer's version counter)
            q = torch.nn.Parameter(torch.empty(0, dtype=full.dtype, device=full.device), requires_grad=True)
            q.data = full[o:o + p.numel()].view(p.shape)
            q._b2_unit = self
            for k, v in vars(p).items():   # carry strategy tags (e.g. tp-replicated) over to the real parameter
                if k.startswith("_b2_tp") or k == "_b2_linear_only":
                    setattr(q, k, v)
            swap[id(p)] = q
            new_params.append(q)
        self.params = new_params
        return swap

    def repoint(self, full):
        self.flat_param = full
        for p, o in zip(self.params, self.offsets):
            p.data = full[o:o + p.numel()].view(p.shape)

    def assign_grads(self, flat_grad):
        self.flat_grad = flat_grad
        if not self.direct:
            flat_grad.zero_()  # autograd accumulates into these views (reference-math path)
        for p, o in zip(self.params, self.offsets):
            p.grad = flat_grad[o:o + p.numel()].view(p.shape)
            p._b2_direct_grad = self.direct
            p._b2_grad_ready = False

    def zero_grad(self):
        if self.flat_grad is None:
            return
        if not self.direct:
            self.flat_grad.zero_()
            return
        for p in self.params:
            p._b2_grad_ready = False

    def finalize_grads(self):
        if not self.direct or self.flat_grad is None:
            return
        for p in self.params:
            if not getattr(p, "_b2_grad_ready", False):
                p.grad.zero_()


class HSDP(FSDP):
    """Hybrid shard: FULL_SHARD inside a replica group (by default one node), replicas kept identical by
    averaging the reduced gradient shards across them.

    Reference: the ``HYBRID_SHARD`` row of the sharding-strategy table
    (``04-fully-sharded-data-parallel/README.md:146-151``; not exposed as a flag there). Ranks
    ``[r * S, (r + 1) * S)`` form replica ``r`` and shard over the peer-memory communicator (NVLink inside
    the node); the ranks holding the same shard index in different replicas all-reduce that shard
    (inter-node traffic is 1/S of a DDP all-reduce). Every replica initialises from the same seed, so
    no parameter broadcast is needed.
    """

    name = "hsdp"
    has_post_reduce = True

    def __init__(self, cfg, device, lr=3e-5, seed=0, dtype=torch.bfloat16, replicas: Optional[int] = None, **kw):
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("HSDP needs an initialised process group")
        world, rank = dist.get_world_size(), dist.get_rank()
        if replicas is None:  # one replica per node
            local = int(os.environ.get("LOCAL_WORLD_SIZE", world))
            replicas = max(world // max(local, 1), 1)
        if world % replicas:
            raise ValueError(f"world size {world} is not a multiple of {replicas} replicas")
        S = world // replicas
        if S < 2:
            raise ValueError("hybrid shard needs at least 2 ranks per replica (use DDP otherwise)")
        self.replicas, self.replica_id = replicas, rank // S
        self.global_rank, self.global_world = rank, world
        # every rank creates every group, in the same order
        shard_groups = [dist.new_group(list(range(r * S, (r + 1) * S))) for r in range(replicas)]
        rep_groups = [dist.new_group(list(range(k, world, S))) for k in range(S)]
        self.rep_group = rep_groups[rank % S] if replicas > 1 else None
        kw.pop("comm", None)
        super().__init__(cfg, device, lr=lr, seed=seed, dtype=dtype, group=shard_groups[rank // S], **kw)

    def _sync_unit(self, full):
        import torch.distributed as dist

        dist.broadcast(full, src=0)   # all replicas, not only this one's sharding group

    def _xchg(self):
        # chunks are cut over the whole world, slices repeat in every replica: the exchange spans all ranks
        S = self.W
        return None, self.global_world, self.global_rank, (lambda r: r % S), (lambda a, b: a // S == b // S)

    def post_reduce(self, i, shard):
        if self.rep_group is None:
            return
        import torch.distributed as dist

        dist.all_reduce(shard, group=self.rep_group)   # SUM: gloo has no AVG
        shard.div_(self.replicas)

    @property
    def rank(self):
        return self.global_rank

    @property
    def world_size(self):
        return self.global_world

    def data_parallel_coords(self):
        return self.global_rank, self.global_world

    def tokens_per_step(self, batch_size, seq_len):
        return self.global_world * batch_size * seq_len

    def barrier(self):
        import torch.distributed as dist

        dist.barrier()
