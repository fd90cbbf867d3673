#!/usr/bin/env python
"""Collective micro-benchmark on N GPUs of one node: peer-memory kernels (csrc/comm.cu) vs NCCL for the
FSDP unit sizes of Llama-3.1-8B, alone and overlapped with a stream of tcgen05 GEMMs.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/comm_bench.py
Times are CUDA events on the launching stream, max over ranks. Roofline: bytes that must cross NVLink
per rank / 770 GB/s (measured peer copy, B200_PROFILING.md).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from b200train.ops.linear import gemm  # noqa: E402
from b200train.parallel.comm import init_distributed  # noqa: E402
from b200train.parallel.peer import PeerComm  # noqa: E402

LINK_GBS = 770.0


def timed(fn, iters=8, warm=3, stream=None):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    rank, local_rank, W, device = init_distributed("cuda")
    comm = PeerComm(None, device)
    numel = 218_112_000 // (8 * W) * (8 * W)      # one Llama-3.1-8B decoder layer
    n = numel // W
    shard = comm.empty(n)
    shard.normal_()
    full_sym = comm.empty(numel)
    full_sym.normal_()
    full = torch.empty(numel, dtype=torch.bfloat16, device=device)
    out_shard = comm.empty(n)
    p, m, v = (torch.zeros(n, dtype=torch.bfloat16, device=device) for _ in range(3))
    res = {"world": W, "unit_mb": numel * 2 / 1e6}
    wire_mb = numel * 2 / 1e6 * (W - 1) / W
    res["roofline_ms"] = wire_mb / 1e3 / LINK_GBS * 1e3

    res["nccl_ag_ms"] = timed(lambda: dist.all_gather_into_tensor(full, shard))
    res["nccl_rs_ms"] = timed(lambda: dist.reduce_scatter_tensor(out_shard, full_sym, op=dist.ReduceOp.AVG))
    res["peer_ag_ce_ms"] = timed(lambda: comm.all_gather(full, shard, copy_engine=True))
    for mb in (32, 64, 128, 148, 296):
        res[f"peer_ag_ms_b{mb}"] = timed(lambda: comm.all_gather(full, shard, max_blocks=mb, copy_engine=False))
        res[f"peer_rs_ms_b{mb}"] = timed(lambda: comm.reduce_scatter(out_shard, full_sym, max_blocks=mb, nvls=False))
        res[f"peer_rs_adamw_ms_b{mb}"] = timed(
            lambda: comm.reduce_scatter_adamw(p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3, max_blocks=mb, nvls=False))

    comm.rs_transport = "ce"
    res["ce_rs_ms"] = timed(lambda: comm.reduce_scatter(out_shard, full_sym))
    for mb in (148, 296):
        res[f"ce_rs_adamw_ms_b{mb}"] = timed(
            lambda: comm.reduce_scatter_adamw(p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3, max_blocks=mb))
    comm.rs_transport = "auto"
    res["nvls"] = bool(comm.has_nvls)
    if comm.has_nvls:
        ar_buf = comm.empty(numel)
        ar_buf.normal_()
        res["nccl_ar_ms"] = timed(lambda: dist.all_reduce(ar_buf, op=dist.ReduceOp.AVG))
        res["peer_ar_ms_b64"] = timed(lambda: comm.all_reduce(ar_buf, max_blocks=64, nvls=False))
        for mb in (16, 32, 64, 148):
            res[f"nvls_rs_ms_b{mb}"] = timed(lambda: comm.reduce_scatter(out_shard, full_sym, max_blocks=mb, nvls=True))
            res[f"nvls_rs_adamw_ms_b{mb}"] = timed(lambda: comm.reduce_scatter_adamw(
                p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3, max_blocks=mb, nvls=True))
            res[f"nvls_ar_ms_b{mb}"] = timed(lambda: comm.all_reduce(ar_buf, max_blocks=mb, nvls=True))

    # overlap: a chain of GEMMs on the main stream with collectives on a side stream
    A = torch.randn(4096, 4096, device=device).to(torch.bfloat16)
    Bm = torch.randn(14336 * 2, 4096, device=device).to(torch.bfloat16)
    C = torch.empty(4096, 14336 * 2, device=device, dtype=torch.bfloat16)
    side = torch.cuda.Stream(device)

    def gemms():
        for _ in range(4):
            gemm(A, Bm, out=C)

    res["gemm_chain_ms"] = timed(gemms)

    def overlapped(coll):
        def fn():
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                coll()
            gemms()
            torch.cuda.current_stream().wait_stream(side)
        return fn

    res["gemm+nccl_ag_rs_ms"] = timed(overlapped(lambda: (dist.all_gather_into_tensor(full, shard),
                                                         dist.reduce_scatter_tensor(out_shard, full_sym,
                                                                                    op=dist.ReduceOp.AVG))))
    res["gemm+peer_agce_rs_ms_b64"] = timed(overlapped(lambda: (comm.all_gather(full, shard, copy_engine=True),
                                                                comm.reduce_scatter(out_shard, full_sym,
                                                                                    max_blocks=64, nvls=False))))
    if comm.has_nvls:
        for mb in (16, 32, 64):
            res[f"gemm+peer_agce_nvls_rs_ms_b{mb}"] = timed(overlapped(
                lambda: (comm.all_gather(full, shard, copy_engine=True),
                         comm.reduce_scatter(out_shard, full_sym, max_blocks=mb, nvls=True))))
    comm.rs_transport = "ce"
    res["gemm+agce_ce_rsadamw_ms"] = timed(overlapped(
        lambda: (comm.all_gather(full, shard, copy_engine=True),
                 comm.reduce_scatter_adamw(p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3))))
    comm.rs_transport = "peer"
    for mb in (32, 148, 296):
        res[f"gemm+peer_agce_rsadamw_ms_b{mb}"] = timed(overlapped(
            lambda: (comm.all_gather(full, shard, copy_engine=True),
                     comm.reduce_scatter_adamw(p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3, max_blocks=mb))))
    for mb in (64, 128):
        res[f"gemm+peer_ag_rs_ms_b{mb}"] = timed(overlapped(lambda: (comm.all_gather(full, shard, max_blocks=mb, copy_engine=False),
                                                                     comm.reduce_scatter(out_shard, full_sym,
                                                                                         max_blocks=mb, nvls=False))))
        res[f"gemm+peer_ag_rsadamw_ms_b{mb}"] = timed(overlapped(
            lambda: (comm.all_gather(full, shard, max_blocks=mb, copy_engine=False),
                     comm.reduce_scatter_adamw(p, m, v, full_sym, 1e-4, 0.9, 0.999, 1e-8, 0.01, 3, max_blocks=mb, nvls=False))))
    if rank == 0:
        for k, val in res.items():
            print(f"{k:36s} {val:10.3f}" if isinstance(val, float) else f"{k:36s} {val}")
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"comm_bench_w{W}.json"), "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
