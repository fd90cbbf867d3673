#!/usr/bin/env python
"""Tiny driver for ``ncu --set full`` captures: launches each hot kernel a few times on Llama-3.1-8B
shapes (pick one with ``-k regex:<name>`` on the ncu command line)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train.ops.attention import attn_bwd_raw, attn_fwd_raw  # noqa: E402
from b200train.ops.linear import gemm  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
T = 4096
if which in ("gemm", "all"):
    A = (torch.randn(T, 4096, device="cuda") * 0.1).to(torch.bfloat16)
    B = (torch.randn(28672, 4096, device="cuda") * 0.1).to(torch.bfloat16)
    C = torch.empty(T, 28672, device="cuda", dtype=torch.bfloat16)
    for _ in range(4):
        gemm(A, B, out=C)
if which in ("attn", "all"):
    nq, nkv, D = 32, 8, 128
    qkv = (torch.randn(T, (nq + 2 * nkv) * D, device="cuda")).to(torch.bfloat16)
    cu = torch.tensor([0, T], dtype=torch.int32, device="cuda")
    q, k, v = qkv[:, :nq * D], qkv[:, nq * D:(nq + nkv) * D], qkv[:, (nq + nkv) * D:]
    dqkv = torch.empty_like(qkv)
    dq, dk, dv = dqkv[:, :nq * D], dqkv[:, nq * D:(nq + nkv) * D], dqkv[:, (nq + nkv) * D:]
    for _ in range(3):
        o, lse = attn_fwd_raw(q, k, v, cu, T, nq, nkv, D, D ** -0.5)
        attn_bwd_raw(torch.randn_like(o), q, k, v, o, lse, dq, dk, dv, cu, T, nq, nkv, D, D ** -0.5)
if which in ("gemm_fused", "all"):
    # the fused-epilogue GEMMs on one Llama-3.1-8B layer: gate/up + SwiGLU, down dgrad + SwiGLU backward, QKV + RoPE
    from b200train import ops

    H, I = 4096, 14336
    x = (torch.randn(T, H, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_()
    w_gu = (torch.randn(2 * I, H, device="cuda") * H ** -0.5).to(torch.bfloat16).requires_grad_()
    w_d = (torch.randn(H, I, device="cuda") * I ** -0.5).to(torch.bfloat16).requires_grad_()
    w_qkv = (torch.randn(6144, H, device="cuda") * H ** -0.5).to(torch.bfloat16)
    cos, sin = ops.rope_tables(128, 8192, 500000.0, None, device="cuda")
    for _ in range(3):
        y = ops.mlp_swiglu(x, w_gu, w_d)
        y.backward(torch.ones_like(y))
        ops.linear_rope(x.detach(), w_qkv, cos, sin, None, 40, 128, T)
torch.cuda.synchronize()
print("ok")
