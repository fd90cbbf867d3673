#!/usr/bin/env python
"""Small-shape pass over every single-GPU kernel for compute-sanitizer (memcheck / racecheck /
synccheck): GEMM in its layouts and epilogues (including the tail-split schedule), attention forward and
both backward kernels, the elementwise / norm / CE / AdamW kernels. Shapes are tiny on purpose: the tool
slows tcgen05 / TMA kernels down by 10-100x.

  compute-sanitizer --tool memcheck python bench/sanitize_driver.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train import ops  # noqa: E402
from b200train.ops.attention import attn_bwd_raw, attn_fwd_raw  # noqa: E402
from b200train.ops.linear import gemm  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"


def rnd(*s):
    return (torch.randn(*s, device="cuda") * 0.3).to(torch.bfloat16)


if which in ("gemm", "all"):
    for M, N, K in [(256, 512, 128), (384, 640, 192)]:
        A, B = rnd(M, K), rnd(N, K)
        gemm(A, B)
        gemm(A, B.t().contiguous(), b_mn=True)
        gemm(A.t().contiguous(), B.t().contiguous(), a_mn=True, b_mn=True)
        C = rnd(M, N)
        gemm(A, B, out=C, epilogue=3)
        gemm(A, B, epilogue=2, bias=rnd(N))
    # tail split: 255 tile pairs on 74 workers
    gemm(rnd(3840, 64), rnd(4160, 64))
    print("gemm ok", flush=True)
if which in ("attn", "all"):
    for D, nq, nkv in [(128, 4, 2), (64, 2, 2)]:
        T = 384
        qkv = rnd(T, (nq + 2 * nkv) * D)
        cu = torch.tensor([0, 200, T], dtype=torch.int32, device="cuda")
        q, k, v = qkv[:, :nq * D], qkv[:, nq * D:(nq + nkv) * D], qkv[:, (nq + nkv) * D:]
        o, lse = attn_fwd_raw(q, k, v, cu, 200, nq, nkv, D, D ** -0.5)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :nq * D], dqkv[:, nq * D:(nq + nkv) * D], dqkv[:, (nq + nkv) * D:]
        attn_bwd_raw(torch.randn_like(o), q, k, v, o, lse, dq, dk, dv, cu, 200, nq, nkv, D, D ** -0.5)
    print("attention ok", flush=True)
if which in ("elementwise", "all"):
    T, H, I = 96, 512, 1024
    x = rnd(T, H).requires_grad_()
    w = rnd(H).requires_grad_()
    y, h = ops.rmsnorm(x, w, 1e-5, rnd(T, H))
    (y.float().sum() + h.float().sum()).backward()
    gu = rnd(T, 2 * I).requires_grad_()
    ops.swiglu(gu).float().sum().backward()
    n = 8 * 1000
    p, g, m, v = (rnd(n) for _ in range(4))
    ops.adamw_step(p, g, m.abs(), v.abs(), 1e-3, step=2)
    from b200train.ops.cross_entropy import linear_cross_entropy

    xe, we = rnd(T, H).requires_grad_(), rnd(1000, H).requires_grad_()
    linear_cross_entropy(xe, we, torch.randint(0, 1000, (T,), device="cuda")).backward()
    emb = rnd(1000, H).requires_grad_()
    ops.embedding(torch.randint(0, 1000, (T,), device="cuda"), emb).float().sum().backward()
    print("elementwise ok", flush=True)
torch.cuda.synchronize()
print("done")
