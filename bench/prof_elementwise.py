#!/usr/bin/env python
"""Driver for ncu captures of the memory-bound kernels (rmsnorm fwd / bwd, swiglu) on Llama-8B shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200train import _C  # noqa: E402

T, H = 4096, 4096
st = torch.cuda.current_stream().cuda_stream
xs = [(torch.randn(T, H, device="cuda") * 0.1).to(torch.bfloat16) for _ in range(4)]
w = torch.ones(H, device="cuda", dtype=torch.bfloat16)
rstd = torch.rand(T, device="cuda") + 0.5
y, h, dx = (torch.empty_like(xs[0]) for _ in range(3))
dw = torch.empty(H, device="cuda", dtype=torch.bfloat16)
partial = torch.empty(_C.load().b2_rmsnorm_bwd_partial_rows(T) * H, device="cuda", dtype=torch.float32)
for i in range(3):
    _C.call("b2_rmsnorm_fwd", y.data_ptr(), rstd.data_ptr(), h.data_ptr(), xs[i].data_ptr(), xs[i + 1].data_ptr(),
            w.data_ptr(), T, H, 1e-5, st)
    _C.call("b2_rmsnorm_bwd", dx.data_ptr(), dw.data_ptr(), partial.data_ptr(), xs[i].data_ptr(), xs[i + 1].data_ptr(),
            w.data_ptr(), rstd.data_ptr(), xs[(i + 2) % 4].data_ptr(), T, H, 0, st)
torch.cuda.synchronize()
print("ok")
