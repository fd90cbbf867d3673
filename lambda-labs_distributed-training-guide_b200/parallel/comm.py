"""Process bootstrap and collectives.

``init_distributed`` mirrors the reference bootstrap (``dist.init_process_group()`` with env://
rendezvous, ``local_rank = rank % device_count``, ``torch.cuda.set_device``:
``02-distributed-data-parallel/train_llm.py:37-54``; MPI launchers supply
``OMPI_COMM_WORLD_RANK/SIZE`` instead: ``03-job-launchers/README.md:127-133``).

``Comm`` is the collective surface the strategies use. ``TorchComm`` issues NCCL (GPU) or gloo (CPU
tests) collectives through ``torch.distributed`` on a side stream -- it is the baseline and the
correctness oracle. ``PeerComm`` (``parallel/peer.py``) provides the same surface with the b200train
NVLink peer-memory kernels and is selected on a single NVSwitch node.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist


def init_distributed(device_type: Optional[str] = None, timeout_s: int = 1800):
    """Initialise the default process group from the launcher's environment. Returns
    ``(rank, local_rank, world_size, device)``; works without a launcher (world 1)."""
    env = os.environ
    # kernels of this framework wait on flags that copy engines and other streams produce (fused gather ->
    # GEMM); streams that share a hardware queue would serialise behind such a waiter, so ask for the
    # maximum number of queues before the CUDA context exists
    env.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    if "RANK" not in env and "OMPI_COMM_WORLD_RANK" in env:  # mpirun
        env["RANK"] = env["OMPI_COMM_WORLD_RANK"]
        env["WORLD_SIZE"] = env["OMPI_COMM_WORLD_SIZE"]
        env.setdefault("LOCAL_RANK", env.get("OMPI_COMM_WORLD_LOCAL_RANK", "0"))
    if device_type is None:
        device_type = "cuda" if torch.cuda.is_available() else "cpu"
    world = int(env.get("WORLD_SIZE", "1"))
    rank = int(env.get("RANK", "0"))
    if device_type == "cuda":
        local_rank = int(env.get("LOCAL_RANK", rank % max(torch.cuda.device_count(), 1)))
        device = torch.device(f"cuda:{local_rank}")
        torch.cuda.set_device(device)
    else:
        local_rank = int(env.get("LOCAL_RANK", "0"))
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        env.setdefault("MASTER_ADDR", "127.0.0.1")
        env.setdefault("MASTER_PORT", "29500")
        backend = "nccl" if device_type == "cuda" else "gloo"
        kw = {}
        if device_type == "cuda":
            kw["device_id"] = device
        dist.init_process_group(backend=backend, timeout=datetime.timedelta(seconds=timeout_s), **kw)
    return rank, local_rank, world, device


class Comm:
    """Collectives over one process group for flat tensors."""

    rank = 0
    world = 1

    def all_gather(self, full: torch.Tensor, shard: torch.Tensor):
        full.copy_(shard)

    def all_gather_inplace(self, full: torch.Tensor):
        """Every rank owns slice ``rank`` of its ``full`` buffer; fill the other slices from their owners."""
        n = full.numel() // self.world
        self.all_gather(full, full[self.rank * n:(self.rank + 1) * n])

    def reduce_scatter(self, shard: torch.Tensor, full: torch.Tensor, average: bool = True):
        shard.copy_(full)

    def all_reduce(self, t: torch.Tensor, average: bool = True):
        return t

    def broadcast(self, t: torch.Tensor, src: int = 0):
        return t

    def barrier(self):
        """Host-blocking barrier (filesystem coordination, start-up)."""

    def stream_barrier(self):
        """Device-side barrier ordered on the current stream (no host blocking). NCCL collectives already
        synchronise the ranks they touch, so the default is a no-op."""

    def empty(self, numel: int, dtype=torch.bfloat16, device=None) -> torch.Tensor:
        """Allocate a buffer that may take part in collectives (symmetric memory for peer comm)."""
        return torch.empty(numel, dtype=dtype, device=device)

    def reduce_scatter_adamw(self, *a, **k) -> bool:
        return False


class TorchComm(Comm):
    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.backend = dist.get_backend(group)

    def all_gather(self, full, shard):
        if self.world == 1:
            if full.data_ptr() != shard.data_ptr():
                full.copy_(shard)
            return
        dist.all_gather_into_tensor(full, shard, group=self.group)

    def reduce_scatter(self, shard, full, average=True):
        if self.world == 1:
            if full.data_ptr() != shard.data_ptr():
                shard.copy_(full)
            return
        if self.backend == "nccl":
            dist.reduce_scatter_tensor(shard, full, op=dist.ReduceOp.AVG if average else dist.ReduceOp.SUM,
                                       group=self.group)
        else:  # gloo has no reduce_scatter: all-reduce then keep the owned slice
            tmp = full.float() if full.dtype == torch.bfloat16 else full.clone()
            dist.all_reduce(tmp, group=self.group)
            if average:
                tmp /= self.world
            n = shard.numel()
            shard.copy_(tmp[self.rank * n:(self.rank + 1) * n])

    def all_reduce(self, t, average=True):
        if self.world == 1:
            return t
        if self.backend == "nccl":
            dist.all_reduce(t, op=dist.ReduceOp.AVG if average else dist.ReduceOp.SUM, group=self.group)
        else:
            tmp = t.float() if t.dtype == torch.bfloat16 else t
            dist.all_reduce(tmp, group=self.group)
            if average:
                tmp /= self.world
            if tmp is not t:
                t.copy_(tmp)
        return t

    def broadcast(self, t, src=0):
        if self.world > 1:
            dist.broadcast(t, src=dist.get_global_rank(self.group, src) if self.group is not None else src,
                           group=self.group)
        return t

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)


def _single_node(group) -> bool:
    """True when every rank of ``group`` runs on this host: peer-mapped memory (and the NVSwitch
    multicast) only spans one NVLink domain; groups that cross nodes use NCCL (or hybrid shard: peer
    memory inside the node, NCCL across)."""
    import socket

    names = [None] * dist.get_world_size(group)
    dist.all_gather_object(names, socket.gethostname(), group=group)
    return len(set(names)) == 1


def make_comm(group=None, prefer_peer: bool = True, device=None) -> Comm:
    if not dist.is_initialized():
        return Comm()
    if prefer_peer and device is not None and torch.device(device).type == "cuda" \
            and os.environ.get("B200TRAIN_COMM", "peer") == "peer" and _single_node(group):
        try:
            from .peer import PeerComm

            return PeerComm(group, device)
        except Exception as e:  # pragma: no cover - depends on the machine
            if os.environ.get("B200TRAIN_COMM") == "peer_strict":
                raise
            import logging

            logging.getLogger(__name__).warning("peer-memory comm unavailable (%s); using NCCL", e)
    return TorchComm(group)
