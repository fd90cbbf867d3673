#!/usr/bin/env python
"""Single-GPU driver for ``ncu`` captures of the LOCAL halves of the fused compute + collective paths (ncu must
not wrap a multi-rank command): the reduce + AdamW kernel of the copy-engine reduce-scatter with 1 and 7 staged
slices (the 2- and 8-rank shapes of one Llama-3.1-8B unit), and the GEMM that waits for gathered chunks, fed by
a flagged gather whose "peers" are other buffers of this GPU."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train import _C  # noqa: E402
from b200train.ops.dispatch import stream  # noqa: E402
from b200train.ops.linear import GatherWait, gemm_gathered  # noqa: E402

dev = torch.device("cuda:0")
unit = 218_112_000 // 64 * 64
for W in (2, 8):
    n = unit // W
    p, m, v, g = (torch.randn(n, device=dev).to(torch.bfloat16) * 0.01 for _ in range(4))
    staging = torch.randn((W - 1) * n, device=dev).to(torch.bfloat16) * 0.01
    for it in range(3):
        _C.call("b2_reduce_adamw", p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), staging.data_ptr(), W - 1, n,
                1.0 / W, 1e-4, 0.9, 0.999, 1e-8, 0.01, it + 1, 148, stream(p))
    torch.cuda.synchronize()
    del p, m, v, g, staging
# flagged gather (4 fake ranks x 4 chunks) feeding the QKV GEMM of a Llama-8B layer
W, sub, rank = 4, 4, 1
N, K, T = 6144, 4096, 4096
n = (N * K + 4096) // (W * sub * 8) * (sub * 8) + sub * 8 * 128
shards = [(torch.randn(n, device=dev) * 0.02).to(torch.bfloat16) for _ in range(W)]
full = torch.zeros(n * W, dtype=torch.bfloat16, device=dev)
flags = torch.zeros(64, dtype=torch.int32, device=dev)
x = (torch.randn(T, K, device=dev) * 0.5).to(torch.bfloat16)
side = torch.cuda.Stream(dev)
arr = (ctypes.c_void_p * W)(*[s.data_ptr() for s in shards])
for epoch in (1, 2, 3):
    with torch.cuda.stream(side):
        _C.call("b2_peer_all_gather_flagged", full.data_ptr(), arr, n * 2, rank, W, sub, 0, 0, flags.data_ptr(), epoch,
                side.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(side)
        _C.call("b2_peer_all_gather_flagged", full.data_ptr(), arr, n * 2, rank, W, sub, 0, 1, flags.data_ptr(), epoch,
                side.cuda_stream)
    torch.cuda.current_stream().wait_event(ev)
    w = full[128:128 + N * K].view(N, K)
    out = gemm_gathered(x, w, GatherWait(flags.data_ptr(), epoch, 128, n // sub, W * sub, sub, rank, W))
    torch.cuda.synchronize()
print("ok")
