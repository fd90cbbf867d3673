"""FSDP all-gather fused into the consuming GEMM (``b2_peer_all_gather_flagged`` + ``b2_gemm_bf16_bwait``).

The gather lands a flat unit in local memory chunk by chunk and raises a flag per chunk; the GEMM whose
weight lives in that unit waits, tile by tile, for the chunks holding the rows it is about to load. The
oracle is the plain sequence on identical inputs -- complete gather, then the ordinary GEMM -- and the
results must be bit-identical (same tiles, same accumulation order; only the order in which tiles are
visited differs). The gathering stream is delayed on purpose so the GEMM really starts before its data.

One-GPU case: the "peers" are other buffers of the same device (the copy / flag / wait machinery is the
same). Multi-GPU case: real peer-mapped shards on every visible GPU (2, 4 or 8).
"""
import ctypes
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _check_fused(srcs_ptrs, shards, rank, W, sub, dev, gather_stream, full_ref, seed):
    from b200train import _C
    from b200train.ops.linear import GatherWait, gemm, gemm_gathered

    n = shards[rank].numel()
    g = torch.Generator(device=dev).manual_seed(seed)
    results = {}
    for name, (M, N, K, b_mn, with_bias) in {
        "fwd": (512, 1536, 256, False, False),
        "fwd_bias": (384, 768, 512, False, True),
        "dgrad": (512, 256, 1536, True, False),          # weight stored [K, N] = [1536, 256]
        "conv1d_fwd": (256, 768, 256, True, True),
    }.items():
        rows, cols = (K, N) if b_mn else (N, K)
        off, numel = 3 * 128, n * W
        assert off + rows * cols <= numel, (name, numel)
        x = (torch.randn(M, K, device=dev, generator=g) * 0.5).to(torch.bfloat16)
        bias = (torch.randn(N, device=dev, generator=g)).to(torch.bfloat16) if with_bias else None
        full = torch.zeros(numel, dtype=torch.bfloat16, device=dev)
        flags = torch.zeros(64, dtype=torch.int32, device=dev)
        for epoch in (1, 2):   # second round: flags carry the previous epoch, the buffer the previous data
            full.zero_()
            torch.cuda.synchronize()
            arr = (ctypes.c_void_p * W)(*srcs_ptrs)
            with torch.cuda.stream(gather_stream):
                _C.call("b2_peer_all_gather_flagged", full.data_ptr(), arr, n * 2, rank, W, sub, 8, 0,
                        flags.data_ptr(), epoch, gather_stream.cuda_stream)
                ev_local = torch.cuda.Event()
                ev_local.record(gather_stream)
                torch.cuda._sleep(3_000_000)   # ~1.5 ms: the GEMM below is running before any peer chunk lands
                _C.call("b2_peer_all_gather_flagged", full.data_ptr(), arr, n * 2, rank, W, sub, 8, 1,
                        flags.data_ptr(), epoch, gather_stream.cuda_stream)
            torch.cuda.current_stream().wait_event(ev_local)
            w = full[off:off + rows * cols].view(rows, cols)
            ag = GatherWait(flags.data_ptr(), epoch, off, n // sub, W * sub, sub, rank, W)
            got = gemm_gathered(x, w, ag, b_mn=b_mn, bias=bias)
            torch.cuda.synchronize()
            assert torch.equal(full, full_ref), f"{name}: gathered buffer differs from the reference gather"
            w_ref = full_ref[off:off + rows * cols].view(rows, cols)
            want = gemm(x, w_ref, b_mn=b_mn, epilogue=1 if with_bias else 0, bias=bias)
            torch.cuda.synchronize()
            results[f"{name}{epoch}"] = bool(torch.equal(got, want))
            ref32 = x.float() @ (w_ref.float() if b_mn else w_ref.float().t())
            if with_bias:
                ref32 += bias.float()
            err = (got.float() - ref32).abs().max().item() / max(ref32.abs().max().item(), 1e-6)
            results[f"{name}{epoch}_fp32"] = err < 2e-2
    return results


def test_gemm_waits_for_gathered_chunks_one_gpu():
    dev = torch.device("cuda:0")
    W, sub, rank = 4, 4, 1
    n = 128 * 8 * sub * 50          # shard elements (multiple of sub * 8)
    g = torch.Generator(device=dev).manual_seed(7)
    shards = [(torch.randn(n, device=dev, generator=g) * 0.05).to(torch.bfloat16) for _ in range(W)]
    full_ref = torch.cat(shards)
    res = _check_fused([s.data_ptr() for s in shards], shards, rank, W, sub, dev, torch.cuda.Stream(dev), full_ref, 11)
    assert all(res.values()), res


def _worker(rank, world, port, ret):
    import torch.distributed as dist

    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device(f"cuda:{rank}")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from b200train.parallel.peer import PeerComm

        comm = PeerComm(None, dev)
        sub = max(1, 16 // world)
        q = 128 * 8 * sub
        n = (819_200 // world + q - 1) // q * q      # shard elements: the unit holds every weight tried below
        shard = comm.empty(n)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        shard.copy_((torch.randn(n, device=dev, generator=g) * 0.05).to(torch.bfloat16))
        full_ref = torch.empty(n * world, dtype=torch.bfloat16, device=dev)
        torch.cuda.synchronize()
        dist.barrier()
        dist.all_gather_into_tensor(full_ref, shard.clone())
        torch.cuda.synchronize()
        ptrs = [int(p) for p in comm._peer_array(shard)]
        res = _check_fused(ptrs, [shard] * world, rank, world, sub, dev, torch.cuda.Stream(dev), full_ref, 5)
        torch.cuda.synchronize()
        dist.barrier()
        ret[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_gemm_waits_for_gathered_chunks_multi_gpu():
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    for r in range(world):
        assert all(ret[r].values()), (r, dict(ret[r]))
