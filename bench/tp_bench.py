#!/usr/bin/env python
"""Tensor-parallel linear micro-benchmark on N GPUs: fused all-gather->GEMM and GEMM->reduce-scatter
kernels vs the NCCL collective + GEMM sequence, Llama-3.1-8B layer shapes, T = 16384 tokens
(the chapter-06 command: batch 16 x seq 1024). Max over ranks, CUDA events.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/tp_bench.py
Roofline per op = max(FLOPs / measured GEMM peak, NVLink bytes / 770 GB/s).
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
from b200train.parallel.tp_fused import FusedTP  # noqa: E402

def _peak_flops():
    """Measured bf16 GEMM peak of this pool (MEASURED_PEAKS.json, written by the driver); the profiling recipe's
    fallback when the file is absent."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) * 1e12
    except Exception:
        return 1590.0e12


PEAK = _peak_flops()
LINK = 770e9


def timed(fn, iters=10, warm=3):
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
    fz = FusedTP(None, rank, W, device)
    T, H, I = 16384, 4096, 14336
    n = T // W
    res = {"world": W, "tokens": T}
    rnd = lambda *s: (torch.randn(*s, device=device) * 0.05).to(torch.bfloat16)  # noqa: E731
    for name, K, N in (("qkv", H, 6144 // W), ("gate_up", H, 2 * I // W)):
        x_local, w = rnd(n, K), rnd(N, K)
        full = torch.empty(T, K, dtype=torch.bfloat16, device=device)

        def unfused():
            dist.all_gather_into_tensor(full, x_local)
            return gemm(full, w)

        def fused():
            buf, ep, _ = fz.gather_rows(x_local)
            return fz.gemm_ag(buf, ep, w)

        a, b = unfused(), fused()
        torch.cuda.synchronize()
        res[f"{name}_maxdiff"] = (a.float() - b.float()).abs().max().item()
        res[f"{name}_colwise_nccl_ms"] = timed(unfused)
        res[f"{name}_colwise_fused_ms"] = timed(fused)
        res[f"{name}_gemm_only_ms"] = timed(lambda: gemm(full, w))
        res[f"{name}_roofline_ms"] = 1e3 * max(2.0 * T * N * K / PEAK, (W - 1) * n * K * 2 / LINK)
    for name, K, N in (("o", H // W, H), ("down", I // W, H)):
        x, w = rnd(T, K), rnd(N, K)
        out = torch.empty(n, N, dtype=torch.bfloat16, device=device)

        def unfused():
            y = gemm(x, w)
            dist.reduce_scatter_tensor(out, y)
            return out

        def fused():
            return fz.gemm_rs(x, w)

        a, b = unfused().clone(), fused()
        torch.cuda.synchronize()
        res[f"{name}_maxdiff"] = (a.float() - b.float()).abs().max().item()
        res[f"{name}_rowwise_nccl_ms"] = timed(unfused)
        res[f"{name}_rowwise_fused_ms"] = timed(fused)
        res[f"{name}_gemm_only_ms"] = timed(lambda: gemm(x, w))
        res[f"{name}_roofline_ms"] = 1e3 * max(2.0 * T * N * K / PEAK, (W - 1) * n * N * 2 / LINK)
    if rank == 0:
        for k, v in res.items():
            print(f"{k:32s} {v:10.4f}" if isinstance(v, float) else f"{k:32s} {v}")
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"tp_bench_w{W}.json"), "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
