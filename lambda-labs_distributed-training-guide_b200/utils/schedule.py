"""Learning-rate schedule of the reference: ``CosineAnnealingLR(optimizer, T_max=1000,
eta_min=lr*1e-2)`` stepped once per optimizer step and periodic beyond ``T_max``
(``01-single-gpu/train_llm.py:68-70``)."""
from __future__ import annotations

import math

import torch

T_MAX = 1000
ETA_MIN_FACTOR = 1e-2


def make_lr_scheduler(optimizer: torch.optim.Optimizer, lr: float):
    return torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=T_MAX, eta_min=lr * ETA_MIN_FACTOR)


def cosine_lr(step: int, lr: float, t_max: int = T_MAX, eta_min_factor: float = ETA_MIN_FACTOR) -> float:
    """Closed form of the schedule above (what the chained torch implementation converges to)."""
    eta_min = lr * eta_min_factor
    return eta_min + (lr - eta_min) * (1 + math.cos(math.pi * step / t_max)) / 2


def scale_lr(base_lr: float, base_batch: int, batch: int, rule: str = "linear") -> float:
    """LR scaling rules from ``related-topics/effective-batch-size-and-lr/README.md:24,37``."""
    k = batch / base_batch
    if rule == "linear":
        return base_lr * k
    if rule == "sqrt":
        return base_lr * math.sqrt(k)
    raise ValueError(rule)
