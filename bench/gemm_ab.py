"""Interleaved A/B timing of two GEMM configs (tail split on / off) so clock and power drift hit both
arms equally: rounds of [A x n, B x n], median per arm."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train.ops.linear import gemm  # noqa: E402

SPLIT, WHOLE = 2 | (256 << 8), 2 | (256 << 8) | (1 << 24)
shapes = [("o_fwd", 4096, 4096, 4096, False, False), ("down_fwd", 4096, 4096, 14336, False, False),
          ("qkv_fwd", 4096, 6144, 4096, False, False), ("gateup_fwd", 4096, 28672, 4096, False, False),
          ("head_fwd", 4096, 128256, 4096, False, False), ("head_wgrad", 128256, 4096, 4096, True, True),
          ("qkv_dgrad", 4096, 4096, 6144, False, True), ("o_wgrad", 4096, 4096, 4096, True, True)]


def rnd(*s):
    return (torch.randn(*s, device="cuda") * 0.1).to(torch.bfloat16)


for name, M, N, K, a_mn, b_mn in shapes:
    n_rot = max(2, int(300e6 // (2 * (M * K + N * K + M * N))) + 1)
    A = [rnd(K, M) if a_mn else rnd(M, K) for _ in range(n_rot)]
    B = [rnd(K, N) if b_mn else rnd(N, K) for _ in range(n_rot)]
    C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    times = {SPLIT: [], WHOLE: []}
    for rnd_i in range(7):
        for cfg in (SPLIT, WHOLE) if rnd_i % 2 == 0 else (WHOLE, SPLIT):
            for i in range(2):
                gemm(A[i % n_rot], B[i % n_rot], a_mn=a_mn, b_mn=b_mn, out=C, config=cfg)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(10):
                gemm(A[i % n_rot], B[i % n_rot], a_mn=a_mn, b_mn=b_mn, out=C, config=cfg)
            e1.record()
            torch.cuda.synchronize()
            times[cfg].append(e0.elapsed_time(e1) / 10)
    med = {k: sorted(v)[len(v) // 2] for k, v in times.items()}
    fl = 2.0 * M * N * K / 1e9
    print(f"{name:12s} split {fl / med[SPLIT]:7.1f} TF  whole {fl / med[WHOLE]:7.1f} TF  ratio {med[WHOLE] / med[SPLIT]:.3f}",
          flush=True)
    del A, B, C
