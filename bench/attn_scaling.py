"""Attention kernel time against sequence length, causal and not, forward and backward: separates the
per-kv-block cost from per-CTA cost and tail effects (fit t * SMs = a * blocks + b * CTAs)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train.ops.attention import attn_bwd_raw, attn_fwd_raw  # noqa: E402

nq, nkv, D = 32, 8, 128


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for S in (1024, 2048, 4096, 8192, 16384):
    for causal in (True, False):
        if S == 16384 and not causal:
            continue
        qkv = (torch.randn(S, (nq + 2 * nkv) * D, device="cuda")).to(torch.bfloat16)
        cu = torch.tensor([0, S], dtype=torch.int32, device="cuda")
        q, k, v = qkv[:, :nq * D], qkv[:, nq * D:(nq + nkv) * D], qkv[:, (nq + nkv) * D:]
        o, lse = attn_fwd_raw(q, k, v, cu, S, nq, nkv, D, D ** -0.5, causal)
        do = torch.randn_like(o)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :nq * D], dqkv[:, nq * D:(nq + nkv) * D], dqkv[:, (nq + nkv) * D:]
        tf = timeit(lambda: attn_fwd_raw(q, k, v, cu, S, nq, nkv, D, D ** -0.5, causal))
        tb = timeit(lambda: attn_bwd_raw(do, q, k, v, o, lse, dq, dk, dv, cu, S, nq, nkv, D, D ** -0.5, causal))
        nb = S // 128
        blocks = nq * (nb * (nb + 1) // 2 if causal else nb * nb)
        fl = 4.0 * blocks * 128 * 128 * D
        print(f"S={S:6d} causal={int(causal)} fwd {tf * 1e3:8.1f} us ({fl / tf / 1e9:6.0f} TF)  bwd {tb * 1e3:8.1f} us "
              f"({2.5 * fl / tb / 1e9:6.0f} TF)  blocks {blocks} ctas_fwd {nq * nb}  us*SM/block fwd {tf * 1e3 * 148 / blocks:.3f}",
              flush=True)
