"""Sharded checkpoints in the ``torch.distributed.checkpoint`` (DCP) directory layout.

The reference writes ``exp_dir/checkpoint/`` (``.metadata`` + ``__<rank>_<n>.distcp``) with keys
``model.<fqn>`` and ``optimizer.*`` from every rank (``04-fully-sharded-data-parallel/train_llm.py:122,
249-263``; Stateful objects in ``06-tensor-parallel/train_llm.py:177-189,283-295``). Here every rank
contributes a dim-0 chunk of each HF-named parameter (and of its two AdamW moments) as a ``DTensor``
over the data-parallel mesh, so the files are readable by stock DCP tooling (``dcp_to_torch_save`` ...)
and by the reference's own ``load`` with the same FQNs. Flat-shard boundaries do not fall on parameter
rows, so saving / loading re-chunks one unit at a time through the strategy's collectives.
"""
from __future__ import annotations

import os
import shutil
from pathlib import Path

import torch
import torch.distributed as dist


def _chunk_bounds(rows: int, world: int, rank: int):
    per = (rows + world - 1) // world
    lo = min(rank * per, rows)
    return lo, min(lo + per, rows)


def _as_dtensor(full: torch.Tensor, mesh, rank: int, world: int):
    """Local dim-0 chunk of ``full`` wrapped as a DTensor of the global shape."""
    from torch.distributed.tensor import DTensor, Shard

    if full.dim() == 0:
        full = full.reshape(1)
    lo, hi = _chunk_bounds(full.shape[0], world, rank)
    local = full[lo:hi].detach().clone().contiguous()
    return DTensor.from_local(local, mesh, [Shard(0)], run_check=False, shape=full.shape,
                              stride=full.contiguous().stride())


def _wrap_chunk(local: torch.Tensor, gshape, mesh):
    """A dim-0 chunk that already is this rank's share of a tensor of shape ``gshape``."""
    from torch.distributed.tensor import DTensor, Shard

    stride = torch.empty(gshape, device="meta").stride()
    return DTensor.from_local(local.contiguous(), mesh, [Shard(0)], run_check=False, shape=torch.Size(gshape),
                              stride=stride)


def _chunk(full: torch.Tensor, mesh, rank: int, world: int):
    if world == 1:
        return full.detach().clone()
    return _as_dtensor(full, mesh, rank, world)


def _chunk_2d(local: torch.Tensor, strategy, name: str):
    """(dp, tp) mesh: ``local`` is this tp rank's slice of the global tensor (already complete over dp).
    tp shards ``tp_dim`` of the global tensor; dp shards a different dim (or dim 0 of replicated 1-D
    tensors) when the strategy shards over dp, else it replicates."""
    from torch.distributed.tensor import DTensor, Replicate, Shard

    m = strategy.mesh
    tp_dim = strategy.model.tp_shard_dim(name)
    if local.dim() == 0:
        local = local.reshape(1)
    gshape = list(local.shape)
    if tp_dim is not None:
        gshape[tp_dim] *= m.tp
    placements = [Replicate(), Shard(tp_dim) if tp_dim is not None else Replicate()]
    piece = local
    if strategy.shards_over_dp and m.dp > 1:
        dp_dim = 0 if tp_dim != 0 else 1
        lo, hi = _chunk_bounds(local.shape[dp_dim], m.dp, m.dp_rank)
        piece = local.narrow(dp_dim, lo, hi - lo)
        placements[0] = Shard(dp_dim)
    piece = piece.detach().clone().contiguous()
    stride = torch.empty(gshape, device="meta").stride()
    return DTensor.from_local(piece, _mesh2d(strategy), placements, run_check=False, shape=torch.Size(gshape),
                              stride=stride)


def _local_from_2d(t, strategy, name: str):
    """Inverse of :func:`_chunk_2d`: this tp rank's slice, complete over dp."""
    from torch.distributed.tensor import DTensor, Replicate, Shard

    if not isinstance(t, DTensor):
        return t
    tp_dim = strategy.model.tp_shard_dim(name)
    return t.redistribute(placements=[Replicate(), Shard(tp_dim) if tp_dim is not None else Replicate()]).to_local()


_MESH2D = {}


def _mesh2d(strategy):
    m = strategy.mesh
    key = (strategy.device.type, m.dp, m.tp)
    if key not in _MESH2D:
        from torch.distributed.device_mesh import init_device_mesh

        _MESH2D[key] = init_device_mesh(strategy.device.type, (m.dp, m.tp), mesh_dim_names=("dp", "tp"))
    return _MESH2D[key]


def build_sharded_state(strategy, mesh):
    """Walk the strategy one unit at a time and keep only this rank's dim-0 chunk of every HF-named
    tensor (parameters and AdamW moments), so the full model never has to exist on one rank."""
    rank, world = strategy.rank, strategy.world_size
    two_d = hasattr(strategy, "mesh") and world > 1
    state = {"model": {}, "optimizer": {"state": {}, "param_groups": strategy.optimizer_param_groups()}}
    exchange = getattr(strategy, "unit_named_chunks", None) if not two_d else None
    for i in range(strategy.num_units()):
        for kind in ("param", "exp_avg", "exp_avg_sq"):
            # flat-sharded strategies hand over their dim-0 chunks after one all-to-all (every element moves at
            # most once); the others expose the whole unit and the chunk is cut out of it
            chunks = exchange(i, kind) if exchange is not None else None
            if chunks is not None:
                pieces = {name: _wrap_chunk(local, gshape, mesh) for name, (local, gshape) in chunks.items()}
            else:
                pieces = {name: _chunk_2d(full, strategy, name) if two_d else _chunk(full, mesh, rank, world)
                          for name, full in strategy.unit_named_full(i, kind).items()}
            for name, piece in pieces.items():
                if kind == "param":
                    state["model"][name] = piece
                else:
                    st = state["optimizer"]["state"].setdefault(
                        name, {"step": torch.tensor(float(strategy.optimizer_step_count()))})
                    st[kind] = piece
    return state


def save_sharded(exp_dir: Path, strategy, state: dict = None, lr_scheduler=None):
    """Every rank calls this (flavour C/D). Rank 0 also writes ``lr_scheduler.pt`` and ``state.json``."""
    from torch.distributed.checkpoint import save

    from .state import save_state

    exp_dir = Path(exp_dir)
    rank = strategy.rank
    strategy.barrier()
    sharded = build_sharded_state(strategy, None if hasattr(strategy, "mesh") else _mesh(strategy))
    # The reference saves into ``checkpoint/`` in place (04-.../train_llm.py:249-263): a crash in the middle
    # leaves a torn directory under an older ``state.json``. Here the new checkpoint is written next to the old
    # one and swapped in by one leader per file system once every rank has finished writing.
    leader = _dir_leader(exp_dir, strategy)
    tmp, final = exp_dir / "checkpoint.tmp", exp_dir / "checkpoint"
    if leader:
        shutil.rmtree(tmp, ignore_errors=True)
    strategy.barrier()
    save(sharded, checkpoint_id=tmp, no_dist=not dist.is_initialized())
    strategy.barrier()
    if leader:
        old = exp_dir / "checkpoint.old"
        shutil.rmtree(old, ignore_errors=True)
        if final.exists():
            os.replace(final, old)
        os.replace(tmp, final)
        shutil.rmtree(old, ignore_errors=True)
    strategy.barrier()
    if rank == 0:
        if lr_scheduler is not None:
            tmp_lr = exp_dir / "lr_scheduler.pt.tmp"
            torch.save(lr_scheduler.state_dict(), tmp_lr)
            os.replace(tmp_lr, exp_dir / "lr_scheduler.pt")
        if state is not None:
            save_state(exp_dir, state)
    strategy.barrier()


def _dir_leader(exp_dir: Path, strategy) -> bool:
    """One rank per distinct file system holding ``exp_dir``: rank 0, plus local rank 0 of every node that
    cannot see a probe file rank 0 has just written (node-local output directories). ``exp_dir.is_mount()``, the
    reference's test (04-.../train_llm.py:158-163), misses shared directories that are not mount points."""
    rank = strategy.rank
    if strategy.world_size == 1:
        exp_dir.mkdir(parents=True, exist_ok=True)
        return True
    probe = exp_dir / ".shared-probe"
    local0 = int(os.environ.get("LOCAL_RANK", "0")) == 0
    if rank == 0 or local0:
        exp_dir.mkdir(parents=True, exist_ok=True)
    if rank == 0:
        probe.write_text("x")
    strategy.barrier()
    leader = rank == 0 or (local0 and not probe.exists())
    strategy.barrier()
    if rank == 0:
        probe.unlink(missing_ok=True)
    return leader


def load_sharded(exp_dir: Path, strategy):
    """Inverse of :func:`save_sharded`: DCP fills the chunk containers in place, then the strategy folds
    them back into its own sharding unit by unit."""
    from torch.distributed.checkpoint import load

    exp_dir = Path(exp_dir)
    two_d = hasattr(strategy, "mesh") and strategy.world_size > 1
    sharded = build_sharded_state(strategy, None if hasattr(strategy, "mesh") else _mesh(strategy))
    load(sharded, checkpoint_id=exp_dir / "checkpoint", no_dist=not dist.is_initialized())
    opt_state = sharded["optimizer"]["state"]
    store = getattr(strategy, "unit_store_chunks", None) if not two_d else None
    for i in range(strategy.num_units()):
        for kind in ("param", "exp_avg", "exp_avg_sq"):
            names = list(strategy.unit_named_full(i, kind, names_only=True))
            if store is not None:
                src = sharded["model"] if kind == "param" else None
                local = {n: _to_local(src[n] if src is not None else opt_state[n][kind]) for n in names}
                if store(i, kind, local):
                    continue
            unpack = (lambda t, n: _local_from_2d(t, strategy, n)) if two_d else (lambda t, n: _to_full(t))
            if kind == "param":
                tensors = {n: unpack(sharded["model"][n], n) for n in names}
            else:
                tensors = {n: unpack(opt_state[n][kind], n) for n in names}
            strategy.unit_store_named(i, kind, tensors)
    steps = [float(_to_full(st["step"])) for st in opt_state.values() if "step" in st]
    strategy.set_optimizer_step_count(int(steps[0]) if steps else 0)
    # the hyper-parameters travel in the checkpoint too (the reference restores them through the DCP
    # optimizer state: 04-fully-sharded-data-parallel/train_llm.py:132-154)
    strategy.set_optimizer_param_groups(sharded["optimizer"]["param_groups"])
    strategy.barrier()


def _to_local(t):
    return t.to_local() if hasattr(t, "to_local") else t


def _to_full(t):
    try:
        from torch.distributed.tensor import DTensor

        if isinstance(t, DTensor):
            return t.full_tensor()
    except ImportError:
        pass
    return t


_MESH = {}


def _mesh(strategy):
    world = strategy.world_size
    if world == 1:
        return None
    key = (strategy.device.type, world)
    if key not in _MESH:
        from torch.distributed.device_mesh import init_device_mesh

        _MESH[key] = init_device_mesh(strategy.device.type, (world,), mesh_dim_names=("dp",))
    return _MESH[key]
