"""AdamW over flat bf16 buffers with bf16 moments (csrc/elementwise.cu ``adamw_kernel``).

Matches ``torch.optim.AdamW(params, lr, fused=True)`` on bf16 parameters (fp32 math per element,
betas (0.9, 0.999), eps 1e-8, weight_decay 0.01): reference ``01-single-gpu/train_llm.py:65``.
"""
from __future__ import annotations

import torch

from .. import _C
from .dispatch import stream, use_kernels


def adamw_step_ref(p, g, m, v, lr, beta1, beta2, eps, wd, step, grad_scale=1.0):
    pf, gf, mf, vf = p.float(), g.float() * grad_scale, m.float(), v.float()
    pf.mul_(1 - lr * wd)
    mf.mul_(beta1).add_(gf, alpha=1 - beta1)
    vf.mul_(beta2).addcmul_(gf, gf, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = vf.sqrt() / (bc2 ** 0.5) + eps
    pf.addcdiv_(mf, denom, value=-lr / bc1)
    p.copy_(pf)
    m.copy_(mf)
    v.copy_(vf)


def adamw_step(p, g, m, v, lr, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.01, step=1, grad_scale=1.0,
               max_blocks=0, p_mc: int = 0):
    """In-place AdamW update of the flat tensors ``p, m, v`` from ``g`` (all same numel/dtype). ``p_mc``: the
    multicast address of ``p`` (ZeRO-1 over symmetric memory): the updated parameters are then stored into
    every rank's replica with ``multimem.st`` instead of being gathered afterwards."""
    assert p.numel() == g.numel() == m.numel() == v.numel()
    if p_mc and use_kernels(p) and p.dtype == torch.bfloat16:
        _C.call("b2_adamw_multicast", p.data_ptr(), p_mc, m.data_ptr(), v.data_ptr(), g.data_ptr(), p.numel(), lr,
                beta1, beta2, eps, wd, step, grad_scale, max_blocks, stream(p))
    elif use_kernels(p) and p.dtype == torch.bfloat16:
        _C.call("b2_adamw", p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), p.numel(), lr, beta1,
                beta2, eps, wd, step, grad_scale, max_blocks, stream(p))
    elif (p.device.type == "cpu" and p.dtype == torch.bfloat16 and _cpu_kernel_ok()
          and all(t.is_contiguous() for t in (p, g, m, v))):
        # offloaded shards: native OpenMP loop over the (pinned) host buffers
        _C.call("b2_cpu_adamw_bf16", p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), p.numel(), lr, beta1,
                beta2, eps, wd, step, grad_scale)
    elif (p.device.type == "cpu" and p.dtype == torch.float32 and grad_scale == 1.0
          and hasattr(torch, "_fused_adamw_")):
        # fp32 on the host (the chapter-01 CPU configuration): ATen's fused CPU AdamW, as the reference's
        # AdamW(fused=True) (01-single-gpu/train_llm.py:65) -- one pass, no temporaries
        with torch.no_grad():
            torch._fused_adamw_([p], [g], [m], [v], [], [torch.tensor(float(step))], lr=lr, beta1=beta1,
                                beta2=beta2, weight_decay=wd, eps=eps, amsgrad=False, maximize=False)
    else:
        with torch.no_grad():
            adamw_step_ref(p, g, m, v, lr, beta1, beta2, eps, wd, step, grad_scale)


_CPU_OK = None


def _cpu_kernel_ok() -> bool:
    """The host AdamW lives in the same library as the CUDA kernels; it is usable wherever that loads."""
    global _CPU_OK
    if _CPU_OK is None:
        try:
            _CPU_OK = hasattr(_C.load(), "b2_cpu_adamw_bf16")
        except Exception:
            _CPU_OK = False
    return _CPU_OK
