"""Tensor parallel (chapter 06) and 2D FSDP x TP (chapter 07) strategies over a (dp, tp) mesh.

Mesh: ``world = dp * tp`` with tp the fastest-varying coordinate (``init_device_mesh("cuda",
(world // tp, tp), ("dp", "tp"))``: ``07-2d-parallel/train_llm.py:48-52``); chapter 06 uses
``(num_nodes, gpus_on_node)`` so TP spans a node and DP spans nodes (``06-.../train_llm.py:41-56``).
All ranks of a tp group read the same batch: the sampler is built over the dp coordinate only
(``06-.../train_llm.py:152-158``) and ``tok/s`` uses ``dp * batch * seq`` (``07-.../train_llm.py:256``).

``TensorParallel`` replicates over dp. The reference forgets to synchronise gradients over dp in chapter
06; here they are all-reduced over dp when dp > 1. ``TwoD`` shards every tp-local unit
over dp with the FSDP machinery (``fully_shard(layer, mesh=mesh["dp"])``: ``07-.../train_llm.py:126-128``).
"""
from __future__ import annotations

from typing import Optional

import os

import torch
import torch.distributed as dist

from ..models import build_model
from ..models.llama_tp import TPLlamaForCausalLM
from ..utils.flat import FlatUnit
from ..utils.schedule import make_lr_scheduler
from .base import Strategy
from .comm import Comm
from .fsdp import FSDP
from .hooks import run_unit
from .optimizer import FlatAdamW
from .tp_ops import TPGroup

_SLICE_INIT_MAX_PARAMS = 2_000_000_000  # below this, initialise by slicing the unsharded model (exact parity)


class Mesh2D:
    """Process groups of a (dp, tp) mesh; every rank creates every group (collective requirement)."""

    def __init__(self, tp: int):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        assert self.world % tp == 0, f"world size {self.world} is not divisible by tp={tp}"
        self.tp, self.dp = tp, self.world // tp
        self.tp_rank, self.dp_rank = self.rank % tp, self.rank // tp
        self.tp_group = self.dp_group = None
        if self.world > 1:
            for d in range(self.dp):
                g = dist.new_group([d * tp + t for t in range(tp)])
                if d == self.dp_rank:
                    self.tp_group = g
            for t in range(tp):
                g = dist.new_group([d * tp + t for d in range(self.dp)])
                if t == self.tp_rank:
                    self.dp_group = g

    def tp_handle(self, device=None, fused: bool = True) -> TPGroup:
        h = TPGroup(self.tp_group, self.tp_rank, self.tp)
        h.fused = None
        if (fused and self.tp > 1 and device is not None and torch.device(device).type == "cuda"
                and os.environ.get("B200TRAIN_TP_FUSED", "1") == "1" and self.tp <= 8):
            from .tp_fused import FusedTP

            h.fused = FusedTP(self.tp_group, self.tp_rank, self.tp, device)
        return h

    def dp_comm(self, device=None) -> Comm:
        """Communicator of this rank's dp group: the peer-memory / NVLS kernels (their own symmetric arenas
        and signal pad per mesh dimension) when the group lives on one node, NCCL otherwise."""
        if self.world == 1 or self.dp == 1:
            c = Comm()
            return c
        from .comm import make_comm

        return make_comm(self.dp_group, True, device)


def _reduce_replicated_grads(units, tp: TPGroup):
    """Sum the partial gradients of tp-replicated parameters (sequence-parallel norm weights) [N11]: every
    such gradient of ``units`` (one unit, or the whole model) goes through ONE peer-memory all-reduce when the
    tp group has a peer handle, else one NCCL call per parameter."""
    if tp.size == 1:
        return
    grads = [p.grad for u in units for p in u.params
             if getattr(p, "_b2_tp_replicated", False) and p.grad is not None]
    if not grads:
        return
    fz = getattr(tp, "fused", None)
    if fz is not None and grads[0].is_cuda and grads[0].dtype == torch.bfloat16:
        fz.all_reduce_sum_(grads)
        return
    for g in grads:
        tp.all_reduce_(g)


def _init_tp_model(model: TPLlamaForCausalLM, cfg, device, dtype, seed):
    if cfg.num_parameters() <= _SLICE_INIT_MAX_PARAMS:
        full = build_model(cfg, dtype=dtype, device=device, seed=seed)
        model.load_from_full(full)
        del full
    else:
        model.init_weights(seed * 7919 + model.tp.rank)


class TensorParallel(Strategy):
    name = "tp"
    shards_over_dp = False  # replicas over dp: checkpoints shard over tp only

    def __init__(self, cfg, device, lr=3e-5, seed=0, dtype=torch.bfloat16, tp: Optional[int] = None,
                 activation_checkpointing: bool = False):
        super().__init__(cfg, device, lr, seed, dtype)
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.mesh = Mesh2D(tp or world)
        self.tpg = self.mesh.tp_handle(self.device)
        self.dp_comm = self.mesh.dp_comm(self.device)
        self.act_ckpt = activation_checkpointing
        self.model = TPLlamaForCausalLM(cfg, self.tpg, dtype=dtype, device=self.device)
        _init_tp_model(self.model, cfg, self.device, dtype, seed)
        self.model.train()
        if self.mesh.dp > 1:
            # gradients are all-reduced over dp: place the flat buffers in the dp communicator's (symmetric) memory
            self.units = [FlatUnit(name, params, pad_to=8 * self.mesh.dp,
                                   alloc=lambda k: self.dp_comm.empty(k, dtype, self.device))
                          for name, params in self.model.units()]
        else:
            self.units = [FlatUnit(name, params) for name, params in self.model.units()]
        self.optimizer = FlatAdamW(self.units, lr=lr)
        self.lr_scheduler = make_lr_scheduler(self.optimizer, lr)
        self.model.unit_runner = self._runner

    def _runner(self, idx, fn, *tensors):
        ckpt = self.act_ckpt and 1 <= idx <= self.cfg.num_hidden_layers
        return run_unit(self, idx, fn, *tensors, activation_checkpoint=ckpt)

    def pre_forward(self, i):
        pass

    def post_forward(self, i):
        pass

    def pre_backward(self, i):
        pass

    def post_backward(self, i):
        u = self.units[i]
        u.finalize_grads()
        if self.mesh.dp > 1:
            # replicas over dp consume the gradient unit by unit: sum the tp-partial norm gradients first
            _reduce_replicated_grads([u], self.tpg)
            self.dp_comm.all_reduce(u.flat_grad, average=True)

    def backward(self, loss):
        loss.backward()
        self.post_backward(0)
        if self.mesh.dp == 1:
            # every sequence-parallel norm-weight gradient of the model in one all-reduce (65 tensors at 8B)
            _reduce_replicated_grads(self.units, self.tpg)

    @property
    def rank(self):
        return self.mesh.rank

    @property
    def world_size(self):
        return self.mesh.world

    def data_parallel_coords(self):
        return self.mesh.dp_rank, self.mesh.dp

    def tokens_per_step(self, batch_size, seq_len):
        return self.mesh.dp * batch_size * seq_len

    def barrier(self):
        if dist.is_initialized():
            dist.barrier()


class TwoD(FSDP):
    """FSDP over the dp coordinate of tp-local units (chapter 07)."""

    name = "2d"
    shards_over_dp = True
    supports_fused_ag = False   # colwise / rowwise already fuse their own (sequence) collectives

    def __init__(self, cfg, device, lr=3e-5, seed=0, dtype=torch.bfloat16, tp: int = 2, **kw):
        self.mesh = Mesh2D(tp)
        self.tpg = self.mesh.tp_handle(device)
        kw.pop("comm", None)
        super().__init__(cfg, device, lr=lr, seed=seed, dtype=dtype, comm=self.mesh.dp_comm(device), **kw)

    def _build_meta_model(self, cfg, dtype):
        return TPLlamaForCausalLM(cfg, self.tpg, dtype=dtype, device="meta")

    def _init_seed(self, seed):
        return seed * 7919 + self.mesh.tp_rank  # distinct shards per tp rank

    def pre_reduce(self, i):
        _reduce_replicated_grads([self.units[i]], self.tpg)

    @property
    def rank(self):
        return self.mesh.rank

    @property
    def world_size(self):
        return self.mesh.world

    def data_parallel_coords(self):
        return self.mesh.dp_rank, self.mesh.dp

    def tokens_per_step(self, batch_size, seq_len):
        return self.mesh.dp * batch_size * seq_len

    def barrier(self):
        if dist.is_initialized():
            dist.barrier()
