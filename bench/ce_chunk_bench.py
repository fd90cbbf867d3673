"""LM head + cross entropy, forward + backward: all rows at once against row chunks (``--ce-chunk-rows``).

Prints one JSON record per (shape, chunk): milliseconds (CUDA events, L2 flushed between iterations, median of
``--iters``) and the peak allocation above the inputs. Shapes: the Llama-3.1-8B head at T=4096 (the benchmark
step) and T=16384 (the chapter-06 shape), the chapter-00 head (V=156,939, H=3072) at T=8192.

    python bench/ce_chunk_bench.py > profiles/ce_chunk_bench.json
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")))


def run(T, H, V, chunk, iters, flush):
    from b200train.ops.cross_entropy import linear_cross_entropy

    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    x = (torch.randn(T, H, device=dev, generator=g) * 0.5).to(torch.bfloat16).requires_grad_()
    w = (torch.randn(V, H, device=dev, generator=g) * 0.02).to(torch.bfloat16).requires_grad_()
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    times = []
    torch.cuda.synchronize()
    base = torch.cuda.memory_allocated()
    peak = 0
    loss = None
    for it in range(iters + 3):
        x.grad = w.grad = None
        flush.zero_()
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = linear_cross_entropy(x, w, labels, chunk_rows=chunk)
        loss.backward()
        e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            times.append(e0.elapsed_time(e1))
            # gradients (x.grad, w.grad) are part of any variant: report what is held beyond inputs + gradients
            peak = max(peak, torch.cuda.max_memory_allocated() - base - (x.numel() + w.numel()) * 2)
    times.sort()
    return {"T": T, "H": H, "V": V, "chunk_rows": chunk, "ms": round(times[len(times) // 2], 4),
            "extra_peak_gb": round(peak / 2 ** 30, 3), "loss": round(loss.item(), 5)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=7)
    a = ap.parse_args()
    flush = torch.empty(256 * 2 ** 20, dtype=torch.uint8, device="cuda")   # larger than the 126 MB L2
    out = []
    for T, H, V, chunks in ((4096, 4096, 128256, (0, 2048, 1024)), (16384, 4096, 128256, (0, 2048)),
                            (8192, 3072, 156939, (0, 2048))):
        for c in chunks:
            rec = run(T, H, V, c, a.iters, flush)
            print(json.dumps(rec), flush=True)
            out.append(rec)
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
