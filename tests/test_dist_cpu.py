"""Multi-process (gloo, world 2) checks of the host-side parallel logic on CPU: DDP (+ZeRO-1), FSDP
(+offload, +activation checkpointing, +no_sync) must reproduce the single-process loss curve on the
same global batch; DCP sharded checkpoints must round-trip; rank-ordering helpers must order."""
import os
import socket
import tempfile

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn_name, args, ret):
    fn = globals()[fn_name]
    _worker_impl(rank, world, port, lambda r, w: fn(r, w, *args), ret)


def _worker_impl(rank, world, port, fn, ret):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


_TRANSPORT_ERRORS = ("connectFullMesh", "Connection refused", "Connection reset by peer", "Address already in use")


def run_dist(fn_name, *args, world=2):
    """Spawn ``world`` gloo ranks. gloo's TCP mesh set-up occasionally fails on a loaded box before any test code
    has run (``connectFullMesh ... Connection refused``): that alone is retried on a fresh port, nothing else is."""
    os.environ.setdefault("GLOO_SOCKET_IFNAME", "lo")
    for attempt in range(3):
        ret = mp.Manager().dict()
        try:
            mp.spawn(_worker, args=(world, _free_port(), fn_name, args, ret), nprocs=world, join=True)
        except Exception as e:
            if attempt < 2 and any(t in str(e) for t in _TRANSPORT_ERRORS):
                continue
            raise
        return [ret[r] for r in range(world)]


def _batches(cfg, n, B, S, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, cfg.vocab_size, (B, S), generator=g) for _ in range(n)]


def _single_losses(strategy, kw, steps=3):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, **kw)
    out = []
    for ids in _batches(cfg, steps, 4, 32):
        out.append(eng.train_step({"input_ids": ids, "labels": ids}).item())
    return out


def w_losses(rank, world, strategy, kw, steps):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, **kw)
    out = []
    for ids in _batches(cfg, steps, 4, 32):
        mine = ids[rank * 2:(rank + 1) * 2]
        loss = eng.train_step({"input_ids": mine, "labels": mine})
        t = loss.detach().clone()
        dist.all_reduce(t)
        out.append((t / world).item())
    return out


def _dist_losses(strategy, kw, steps=3):
    return run_dist("w_losses", strategy, kw, steps)[0]


@pytest.mark.parametrize("strategy,kw", [
    ("ddp", dict(zero1=False)),
    ("ddp", dict(zero1=True)),
    ("fsdp", dict()),
    ("fsdp", dict(cpu_offload=True)),
    ("fsdp", dict(activation_checkpointing=True, reshard_after_forward=False)),
])
def test_matches_single_process(strategy, kw):
    ref = _single_losses(strategy, {k: v for k, v in kw.items()})
    got = _dist_losses(strategy, kw)
    for a, b in zip(ref, got):
        assert abs(a - b) < 2e-3, (ref, got)


def w_no_sync(rank, world):
    if True:
        from b200train.models import load_config
        from b200train.parallel import make_strategy

        cfg = load_config("llama-tiny")
        a = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32)
        b = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32)
        ids = _batches(cfg, 1, 8, 32)[0][rank * 4:(rank + 1) * 4]
        # a: one step on 4 rows; b: two micro-steps of 2 rows with no_sync on the first
        a.train_step({"input_ids": ids, "labels": ids})
        b.zero_grad()
        with b.no_sync():
            out = b.forward({"input_ids": ids[:2], "labels": ids[:2]})
            b.backward(out.loss / 2)
        out = b.forward({"input_ids": ids[2:], "labels": ids[2:]})
        b.backward(out.loss / 2)
        b.optimizer_step()
        la = a.train_step({"input_ids": ids, "labels": ids}).item()
        lb = b.train_step({"input_ids": ids, "labels": ids}).item()
        return la, lb


def test_fsdp_no_sync_accumulation():
    la, lb = run_dist("w_no_sync")[0]
    assert abs(la - lb) < 2e-3, (la, lb)


def w_fused_step_count(rank, world):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    out = []
    for fuse in (True, "kernel", "stream"):
        # the gloo communicator has no reduce-scatter + AdamW kernel: the request falls back to a separate update
        eng = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, fuse_optimizer=fuse)
        ids = _batches(cfg, 1, 4, 32)[0][rank * 2:(rank + 1) * 2]
        for _ in range(3):
            eng.train_step({"input_ids": ids, "labels": ids})
        out.append(eng.optimizer._step)
    return out


def test_fused_optimizer_fallback_counts_each_step_once():
    """`fuse_optimizer` on a communicator without the fused kernel must not advance AdamW's step counter
    twice per step (bias correction and the checkpointed `step` depend on it)."""
    assert run_dist("w_fused_step_count")[0] == [3, 3, 3]


def w_dcp(rank, world, tmp):
    if True:
        from pathlib import Path

        from b200train.ckpt.dcp import load_sharded, save_sharded
        from b200train.models import load_config
        from b200train.parallel import make_strategy

        cfg = load_config("llama-tiny")
        eng = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32)
        ids = _batches(cfg, 1, 4, 32)[0][rank * 2:(rank + 1) * 2]
        for _ in range(2):
            eng.train_step({"input_ids": ids, "labels": ids})
        save_sharded(Path(tmp), eng, state={"epoch": 0, "global_step": 2, "epoch_step": 2, "running_loss": 0.0},
                     lr_scheduler=eng.lr_scheduler)
        nxt = eng.train_step({"input_ids": ids, "labels": ids}).item()
        eng2 = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=7)
        load_sharded(Path(tmp), eng2)
        eng2.lr_scheduler.load_state_dict(torch.load(Path(tmp) / "lr_scheduler.pt"))
        got = eng2.train_step({"input_ids": ids, "labels": ids}).item()
        files = sorted(p.name for p in (Path(tmp) / "checkpoint").iterdir())
        return nxt, got, files


def w_dcp_swap(rank, world, tmp):
    from pathlib import Path

    from b200train.ckpt.dcp import load_sharded, save_sharded
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32)
    ids = _batches(cfg, 1, 4, 32)[0][rank * 2:(rank + 1) * 2]
    eng.train_step({"input_ids": ids, "labels": ids})
    save_sharded(Path(tmp), eng, state={"epoch": 0, "global_step": 1, "epoch_step": 1, "running_loss": 0.0})
    eng.train_step({"input_ids": ids, "labels": ids})
    if rank == 0:   # what a crash in the middle of the next save leaves behind
        (Path(tmp) / "checkpoint.tmp").mkdir()
        (Path(tmp) / "checkpoint.tmp" / "__0_0.distcp").write_bytes(b"torn")
    dist.barrier()
    # the first checkpoint is still whole under the old state.json: a resume at this point works
    probe = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=5)
    load_sharded(Path(tmp), probe)
    first_step = probe.optimizer_step_count()
    save_sharded(Path(tmp), eng, state={"epoch": 0, "global_step": 2, "epoch_step": 2, "running_loss": 0.0})
    nxt = eng.train_step({"input_ids": ids, "labels": ids}).item()
    eng2 = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=7)
    load_sharded(Path(tmp), eng2)
    got = eng2.train_step({"input_ids": ids, "labels": ids}).item()
    return first_step, eng2.optimizer_step_count(), nxt, got, sorted(p.name for p in Path(tmp).iterdir())


def test_dcp_save_swaps_directories_and_survives_a_torn_attempt():
    """A sharded save goes to ``checkpoint.tmp`` and replaces ``checkpoint`` only when every rank has written: the
    previous checkpoint stays loadable until then, leftovers of a crashed attempt are discarded, and no temporary
    directory remains afterwards."""
    tmp = tempfile.mkdtemp()
    first_step, last_step, nxt, got, names = run_dist("w_dcp_swap", tmp)[0]
    assert first_step == 1 and last_step == 3, (first_step, last_step)   # 2 saved + the step taken after loading
    assert abs(nxt - got) < 1e-5, (nxt, got)
    assert "checkpoint" in names and "state.json" in names, names
    assert not any(n.startswith("checkpoint.") or n.startswith(".shared") for n in names), names


def _moment_energy(eng):
    """Sum of squares of all first moments over all ranks (a fingerprint of the optimizer state)."""
    t = torch.zeros(1, dtype=torch.float64)
    for m in eng.optimizer.exp_avg:
        t += m.double().pow(2).sum()
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t)
    return t.item()


def _ws_engine(strategy, seed=0):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    kw = dict(zero1=True) if strategy == "ddp" else {}
    return cfg, make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=seed, **kw)


def w_ws_save(rank, world, tmp, strategy="fsdp"):
    from pathlib import Path

    from b200train.ckpt.dcp import save_sharded

    cfg, eng = _ws_engine(strategy)
    ids = _batches(cfg, 1, 4, 32)[0]
    for _ in range(2):
        eng.train_step({"input_ids": ids[rank::world], "labels": ids[rank::world]})
    save_sharded(Path(tmp), eng, state={"epoch": 0, "global_step": 2, "epoch_step": 2, "running_loss": 0.0})
    with torch.no_grad():
        probe = eng.forward(eng.to_device({"input_ids": ids, "labels": ids})).loss.item()
    return probe, _moment_energy(eng)


def w_ws_load(rank, world, tmp, strategy="fsdp"):
    from pathlib import Path

    from b200train.ckpt.dcp import load_sharded

    cfg, eng = _ws_engine(strategy, seed=99)
    load_sharded(Path(tmp), eng)
    ids = _batches(cfg, 1, 4, 32)[0]
    with torch.no_grad():
        probe = eng.forward(eng.to_device({"input_ids": ids, "labels": ids})).loss.item()
    return probe, _moment_energy(eng), eng.optimizer_step_count()


@pytest.mark.parametrize("strategy,new_worlds", [("fsdp", (3, 1)), ("ddp", (3,))])
def test_sharded_checkpoint_loads_into_a_different_world_size(strategy, new_worlds):
    """Elastic restarts change the number of ranks: a checkpoint written by 2 ranks must load into 3 (chunks that
    do not divide evenly) and into 1 (nothing sharded), with the same weights, moments and step count -- for FSDP
    and for DDP with a sharded (ZeRO-1) optimizer."""
    tmp = tempfile.mkdtemp()
    probe, energy = run_dist("w_ws_save", tmp, strategy, world=2)[0]
    for new_world in new_worlds:
        got = run_dist("w_ws_load", tmp, strategy, world=new_world)[0]
        assert abs(got[0] - probe) < 1e-5, (new_world, got, probe)
        assert abs(got[1] - energy) <= 1e-6 * max(energy, 1e-12), (new_world, got, energy)
        assert got[2] == 2


def w_dcp_stock(rank, world, tmp):
    """A checkpoint directory written by STOCK torch.distributed.checkpoint code -- DTensors sharded on dim 0
    under the reference's keys (``model.<hf name>``, ``optimizer.state.<hf name>.exp_avg`` ...), which is
    what ``04-fully-sharded-data-parallel/train_llm.py:252-258`` produces -- must load into this framework's
    flat-shard FSDP, and a directory written here must load back into plain DTensor containers."""
    from pathlib import Path

    import torch.distributed.checkpoint as DCP
    from torch.distributed.device_mesh import init_device_mesh
    from torch.distributed.tensor import Shard, distribute_tensor

    from b200train.ckpt.dcp import load_sharded, save_sharded
    from b200train.models import build_model, load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    mesh = init_device_mesh("cpu", (world,), mesh_dim_names=("dp",))
    src = build_model(cfg, device="cpu", dtype=torch.float32, seed=123)       # the "other framework's" weights
    sd = {k: v.detach().clone() for k, v in src.hf_state_dict().items()}
    state = {"model": {}, "optimizer": {"state": {}, "param_groups": [
        {"lr": 5e-4, "betas": [0.9, 0.999], "eps": 1e-8, "weight_decay": 0.01, "initial_lr": 1e-3}]}}
    for k, v in sd.items():
        v2 = v.reshape(1) if v.dim() == 0 else v
        state["model"][k] = distribute_tensor(v2, mesh, [Shard(0)])
        state["optimizer"]["state"][k] = {"step": torch.tensor(7.0),
                                          "exp_avg": distribute_tensor(v2 * 0.5, mesh, [Shard(0)]),
                                          "exp_avg_sq": distribute_tensor(v2 * v2, mesh, [Shard(0)])}
    DCP.save(state, checkpoint_id=Path(tmp) / "stock" / "checkpoint")
    eng = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=5)
    load_sharded(Path(tmp) / "stock", eng)
    got = eng.model_state_dict()
    worst = max((got[k].float() - sd[k].float()).abs().max().item() for k in sd)
    lr, step = eng.optimizer.param_groups[0]["lr"], eng.optimizer._step
    # the other direction: our directory into plain DTensor containers
    save_sharded(Path(tmp) / "ours", eng, state={"epoch": 0, "global_step": 7, "epoch_step": 7, "running_loss": 0.0},
                 lr_scheduler=eng.lr_scheduler)
    back = {"model": {k: distribute_tensor(torch.zeros_like(v.reshape(1) if v.dim() == 0 else v), mesh, [Shard(0)])
                      for k, v in sd.items()}}
    DCP.load(back, checkpoint_id=Path(tmp) / "ours" / "checkpoint")
    worst_back = max((back["model"][k].full_tensor().reshape(sd[k].shape) - sd[k]).abs().max().item() for k in sd)
    return worst, lr, step, worst_back


def test_dcp_interoperates_with_stock_dtensor_checkpoints():
    tmp = tempfile.mkdtemp()
    worst, lr, step, worst_back = run_dist("w_dcp_stock", tmp)[0]
    assert worst == 0.0 and worst_back == 0.0, (worst, worst_back)
    assert abs(lr - 5e-4) < 1e-12 and step == 7, (lr, step)


def test_dcp_sharded_roundtrip():
    tmp = tempfile.mkdtemp()
    nxt, got, files = run_dist("w_dcp", tmp)[0]
    assert abs(nxt - got) < 1e-5, (nxt, got)
    assert ".metadata" in files and "__0_0.distcp" in files and "__1_0.distcp" in files, files
    assert os.path.exists(os.path.join(tmp, "state.json"))


def w_chunk_exchange(rank, world, strategy="fsdp"):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    kw = dict(replicas=2) if strategy == "hsdp" else {}
    eng = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=3, **kw)
    ids = _batches(cfg, 1, 2, 32, seed=rank)[0]
    eng.train_step({"input_ids": ids, "labels": ids})       # non-trivial moments
    worst, n_tensors, moved = 0.0, 0, 0
    for i in range(eng.num_units()):
        for kind in ("param", "exp_avg", "exp_avg_sq"):
            chunks = eng.unit_named_chunks(i, kind)
            assert chunks is not None
            full = {k: v.clone() for k, v in eng.unit_named_full(i, kind).items()}
            assert set(chunks) == set(full)
            for name, (local, gshape) in chunks.items():
                want = full[name].reshape(gshape)
                per = (gshape[0] + world - 1) // world
                lo = min(rank * per, gshape[0])
                want = want[lo:min(lo + per, gshape[0])]
                assert local.shape == want.shape, (name, local.shape, want.shape)
                worst = max(worst, (local - want).abs().max().item() if want.numel() else 0.0)
                n_tensors += 1
                moved += local.numel()
            # the inverse exchange: store 2 x the chunks, the gathered unit must come back doubled
            assert eng.unit_store_chunks(i, kind, {k: v[0] * 2 for k, v in chunks.items()})
            back = eng.unit_named_full(i, kind)
            worst = max(worst, max((back[k] - 2 * full[k]).abs().max().item() for k in full))
    return worst, n_tensors, moved


def test_hybrid_shard_checkpoint_chunks_come_from_one_exchange():
    """2 replicas x 2 shards: chunks are cut over all 4 ranks, each rank is served by its own replica; on load every
    replica receives every chunk."""
    for worst, n_tensors, _ in run_dist("w_chunk_exchange", "hsdp", world=4):
        assert worst == 0.0 and n_tensors > 10, (worst, n_tensors)


@pytest.mark.parametrize("world", [3])
def test_fsdp_checkpoint_chunks_come_from_one_exchange(world):
    """DCP's dim-0 chunks are produced from the flat shards by an all-to-all (no full-size gather): they equal the
    chunks cut from the gathered unit for every tensor and moment, also when the rows do not divide by the world
    size, and the inverse exchange puts every element back into the right flat slice."""
    res = run_dist("w_chunk_exchange", world=world)
    total = sum(r[2] for r in res)
    for worst, n_tensors, _ in res:
        assert worst == 0.0 and n_tensors > 10, (worst, n_tensors)
    assert total > 0


def w_sync_states(rank, world, sync):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy("fsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, seed=100 + rank, sync_module_states=sync)
    sd = eng.gather_full_state_dict()
    return {k: v.clone() for k, v in sd.items()}


def test_fsdp_sync_module_states_broadcasts_rank0_init():
    """Ranks seeded differently: with ``sync_module_states`` the sharded model is rank 0's initialisation (N1,
    rank-0 init + broadcast); without it the shards are cut from different initialisations."""
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    synced = run_dist("w_sync_states", True)
    ref = make_strategy("fsdp", load_config("llama-tiny"), "cpu", lr=1e-3, dtype=torch.float32,
                        seed=100).gather_full_state_dict()
    for k, v in ref.items():
        assert torch.equal(synced[0][k], v) and torch.equal(synced[1][k], v), k
    loose = run_dist("w_sync_states", False)
    assert any(not torch.equal(loose[0][k], v) for k, v in ref.items())


def w_order(rank, world, tmp):
    if True:
        import time

        from b200train.cli.train_llm import rank0_first

        with rank0_first():
            if rank == 0:
                time.sleep(0.3)
            with open(os.path.join(tmp, "order.txt"), "a") as f:
                f.write(f"{rank}\n")
        return True


def test_rank_ordered_runs_first_rank_first():
    tmp = tempfile.mkdtemp()
    run_dist("w_order", tmp)
    assert open(os.path.join(tmp, "order.txt")).read().split()[0] == "0"


def w_tp_losses(rank, world, strategy, tp, steps, repeat=False):
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, tp=tp)
    dp_rank, dp = eng.data_parallel_coords()
    out = []
    data = _batches(cfg, steps, 4, 32)
    for ids in ([data[0]] * steps if repeat else data):
        per = 4 // dp
        mine = ids[dp_rank * per:(dp_rank + 1) * per]
        loss = eng.train_step({"input_ids": mine, "labels": mine})
        t = loss.detach().clone()
        dist.all_reduce(t)
        out.append((t / world).item())
    return out


def test_tensor_parallel_matches_single_device():
    """TP=2 with sequence + loss parallel reproduces the unsharded model's loss curve."""
    ref = _single_losses("single", {})
    got = run_dist("w_tp_losses", "tp", 2, 3)[0]
    for a, b in zip(ref, got):
        assert abs(a - b) < 2e-3, (ref, got)


def test_2d_runs_and_learns():
    """2D (dp=2 x tp=2 on 4 processes): FSDP over dp of tp-local units; loss must fall."""
    got = run_dist("w_tp_losses", "2d", 2, 4, True, world=4)[0]
    assert got[-1] < got[0] - 0.3, got


def w_dcp_2d(rank, world, tmp, strategy):
    from pathlib import Path

    from b200train.ckpt.dcp import load_sharded, save_sharded
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, tp=2)
    dp_rank, dp = eng.data_parallel_coords()
    ids = _batches(cfg, 1, 4, 32)[0]
    per = 4 // dp
    mine = ids[dp_rank * per:(dp_rank + 1) * per]
    for _ in range(2):
        eng.train_step({"input_ids": mine, "labels": mine})
    save_sharded(Path(tmp), eng, state={"epoch": 0, "global_step": 2, "epoch_step": 2, "running_loss": 0.0},
                 lr_scheduler=eng.lr_scheduler)
    nxt = eng.train_step({"input_ids": mine, "labels": mine}).item()
    eng2 = make_strategy(strategy, cfg, "cpu", lr=1e-3, dtype=torch.float32, tp=2, seed=11)
    load_sharded(Path(tmp), eng2)
    eng2.lr_scheduler.load_state_dict(torch.load(Path(tmp) / "lr_scheduler.pt"))
    got = eng2.train_step({"input_ids": mine, "labels": mine}).item()
    return nxt, got


@pytest.mark.parametrize("strategy,world", [("tp", 2), ("2d", 4)])
def test_dcp_roundtrip_tensor_parallel(strategy, world):
    tmp = tempfile.mkdtemp()
    nxt, got = run_dist("w_dcp_2d", tmp, strategy, world=world)[0]
    assert abs(nxt - got) < 1e-5, (nxt, got)


def test_fsdp_size_based_wrap_groups_layers():
    """--numel-to-wrap merges consecutive layers into one sharding unit without changing the maths."""
    from b200train.models import build_model, load_config
    from b200train.models.config import group_by_numel

    assert group_by_numel([5, 5, 5, 5, 5], 10) == [[0, 1], [2, 3, 4]]
    assert group_by_numel([20, 3, 20], 10) == [[0], [1, 2]]
    assert group_by_numel([1, 1], 100) == [[0, 1]]
    cfg = load_config("llama-tiny")
    model = build_model(cfg, dtype=torch.float32, device="meta", init=False)
    per_layer = sum(p.numel() for p in model.model.layers[0].parameters())
    n_before = len(model.units())
    model.group_layers(2 * per_layer)
    assert len(model.units()) == n_before - cfg.num_hidden_layers + cfg.num_hidden_layers // 2
    kw = dict(numel_to_wrap=2 * per_layer)  # weights are initialised per unit: same layout on both sides
    ref = _single_losses("fsdp", kw)
    got = _dist_losses("fsdp", kw)
    for a, b in zip(ref, got):
        assert abs(a - b) < 2e-3, (ref, got)


def w_hsdp(rank, world, tmp):
    """2 replicas x 2 shards: losses of 3 steps, then a DCP round trip into a fresh engine."""
    from b200train.ckpt.dcp import load_sharded, save_sharded
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    eng = make_strategy("hsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, replicas=2)
    assert eng.world_size == 4 and eng.W == 2 and eng.tokens_per_step(1, 32) == 4 * 32
    out = []
    for ids in _batches(cfg, 3, 4, 32):
        mine = ids[rank:rank + 1]
        t = eng.train_step({"input_ids": mine, "labels": mine}).detach().clone()
        dist.all_reduce(t)
        out.append((t / world).item())
    # replicas hold identical shards after training
    same = [torch.zeros_like(eng.shard_param[1]) for _ in range(2)]
    dist.all_gather(same, eng.shard_param[1], group=eng.rep_group)
    assert torch.equal(same[0], same[1])
    save_sharded(tmp, eng)
    eng2 = make_strategy("hsdp", cfg, "cpu", lr=1e-3, dtype=torch.float32, replicas=2, seed=123)
    load_sharded(tmp, eng2)
    for a, b in zip(eng.shard_param, eng2.shard_param):
        assert torch.equal(a, b)
    ids = _batches(cfg, 1, 4, 32, seed=9)[0][rank:rank + 1]
    la = eng.train_step({"input_ids": ids, "labels": ids}).item()
    lb = eng2.train_step({"input_ids": ids, "labels": ids}).item()
    assert abs(la - lb) < 1e-5, (la, lb)
    return out


def test_hybrid_shard_matches_single_process(tmp_path):
    ref = _single_losses("fsdp", {})
    got = run_dist("w_hsdp", str(tmp_path), world=4)[0]
    for a, b in zip(ref, got):
        assert abs(a - b) < 2e-3, (ref, got)


def w_hsdp_variants(rank, world):
    """Hybrid shard: CPU offload, activation checkpointing and no_sync accumulation follow the same curve."""
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    ids = _batches(cfg, 1, 4, 32)[0][rank:rank + 1]
    out = {}
    for name, kw in (("plain", {}), ("offload", dict(cpu_offload=True)), ("ckpt", dict(activation_checkpointing=True)),
                     ("accum", {})):
        eng = make_strategy("hsdp", cfg, "cpu", lr=1e-2, dtype=torch.float32, replicas=2, **kw)
        ls = []
        for _ in range(3):
            if name == "accum":   # two identical half-weight micro-batches == one step
                with eng.no_sync():
                    eng.backward(eng.forward({"input_ids": ids, "labels": ids}).loss / 2)
                o = eng.forward({"input_ids": ids, "labels": ids})
                eng.backward(o.loss / 2)
                eng.optimizer_step()
                eng.zero_grad()
                ls.append(o.loss.item())
            else:
                ls.append(eng.train_step({"input_ids": ids, "labels": ids}).item())
        out[name] = ls
    return out


def test_hybrid_shard_variants_agree():
    got = run_dist("w_hsdp_variants", world=4)[0]
    for name in ("offload", "ckpt", "accum"):
        for a, b in zip(got["plain"], got[name]):
            assert abs(a - b) < 1e-4, got
