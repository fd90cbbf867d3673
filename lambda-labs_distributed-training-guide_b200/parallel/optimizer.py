"""AdamW on flat bf16 storage, exposed as a ``torch.optim.Optimizer`` so stock LR schedulers and
``state_dict()`` consumers keep working.

Reference: ``torch.optim.AdamW(model.parameters(), lr=args.lr, fused=True)``
(``01-single-gpu/train_llm.py:65``) -- bf16 parameters, bf16 moments, no master weights, no clipping;
ZeRO-1 sharding of the moments: ``ZeroRedundancyOptimizer`` (``02-distributed-data-parallel/train_llm.py:81-83``).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..ops.optim import adamw_step
from ..utils.flat import FlatUnit


class FlatAdamW(torch.optim.Optimizer):
    """One fused kernel launch per :class:`FlatUnit` (or per owned shard of it).

    ``shard = (rank, world)`` restricts state and updates to this rank's contiguous 1/world slice of
    every unit (ZeRO-1 / FSDP); the caller is responsible for re-assembling parameters afterwards.
    """

    def __init__(self, units: List[FlatUnit], lr=3e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01,
                 shard: Optional[tuple] = None, grad_scale: float = 1.0, moments_device=None, pin: bool = False):
        self.units = units
        self.shard = shard
        self.grad_scale = grad_scale
        params = [p for u in units for p in u.params]
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._step = 0
        self.exp_avg, self.exp_avg_sq = [], []
        for u in units:
            lo, hi = self.unit_range(u)
            dev = moments_device if moments_device is not None else u.device
            kw = dict(pin_memory=True) if (pin and str(dev) == "cpu") else {}
            self.exp_avg.append(torch.zeros(hi - lo, dtype=u.dtype, device=dev, **kw))
            self.exp_avg_sq.append(torch.zeros(hi - lo, dtype=u.dtype, device=dev, **kw))

    def unit_range(self, u: FlatUnit):
        if self.shard is None:
            return 0, u.numel
        rank, world = self.shard
        n = u.numel // world
        assert n * world == u.numel, "unit size must be padded to a multiple of the shard count"
        return rank * n, (rank + 1) * n

    # ---- stepping -------------------------------------------------------------------------------
    def begin_step(self):
        """Advance the step counter once per optimizer step (bias correction is shared by all units)."""
        self._step += 1
        self._opt_called = True  # lets torch LR schedulers see that an optimizer step happened

    def step_unit(self, i: int, max_blocks: int = 0, grad: Optional[torch.Tensor] = None,
                  param: Optional[torch.Tensor] = None, p_mc: int = 0):
        """Update unit ``i``. ``grad`` / ``param`` override the flat buffers (sharded strategies pass
        their reduced gradient shard and their owned parameter shard)."""
        u = self.units[i]
        g = self.param_groups[0]
        lo, hi = self.unit_range(u)
        p = param if param is not None else u.flat_param[lo:hi]
        gr = grad if grad is not None else u.flat_grad[lo:hi]
        adamw_step(p, gr, self.exp_avg[i], self.exp_avg_sq[i], g["lr"], g["betas"][0], g["betas"][1], g["eps"],
                   g["weight_decay"], self._step, self.grad_scale, max_blocks, p_mc=p_mc)

    @torch.no_grad()
    def step(self, closure=None):
        self.begin_step()
        for i, u in enumerate(self.units):
            u.finalize_grads()
            self.step_unit(i)

    def zero_grad(self, set_to_none: bool = True):
        # gradients live in the flat buffers and are overwritten by the next backward
        for u in self.units:
            u.zero_grad()

    # ---- state dict in the torch.optim layout (per-parameter views of the flat moments) ---------
    def state_dict(self):
        state = {}
        idx = 0
        for ui, u in enumerate(self.units):
            lo, hi = self.unit_range(u)
            for p, off, n in u.named_slices():
                s, e = max(off, lo), min(off + n, hi)
                if s < e:
                    state[idx] = {
                        "step": torch.tensor(float(self._step)),
                        "exp_avg": self.exp_avg[ui][s - lo:e - lo],
                        "exp_avg_sq": self.exp_avg_sq[ui][s - lo:e - lo],
                        "shard_range": (s - off, e - off),
                    }
                idx += 1
        g = dict(self.param_groups[0])
        g["params"] = list(range(idx))
        return {"state": state, "param_groups": [g], "step": self._step}

    def load_state_dict(self, sd):
        self._step = int(sd.get("step", 0))
        idx = 0
        for ui, u in enumerate(self.units):
            lo, hi = self.unit_range(u)
            for p, off, n in u.named_slices():
                s, e = max(off, lo), min(off + n, hi)
                st = sd["state"].get(idx, sd["state"].get(str(idx)))
                if st is not None and s < e:
                    self.exp_avg[ui][s - lo:e - lo].copy_(st["exp_avg"].reshape(-1))
                    self.exp_avg_sq[ui][s - lo:e - lo].copy_(st["exp_avg_sq"].reshape(-1))
                    if not self._step:
                        self._step = int(float(st["step"]))
                idx += 1
        g = sd["param_groups"][0]
        for k in ("lr", "betas", "eps", "weight_decay", "initial_lr"):
            if k in g:
                self.param_groups[0][k] = g[k] if k != "betas" else tuple(g[k])
