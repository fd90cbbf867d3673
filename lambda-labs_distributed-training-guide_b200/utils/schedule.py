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


def warmup_cosine_ratio(step: int, total_num_steps: int, warmup_num_steps: int = 1000, cos_min_ratio: float = 1e-4,
                        warmup_min_ratio: float = 0.0) -> float:
    """Multiplier of DeepSpeed's ``WarmupCosineLR`` (the schedule of the reference's ``ds_config.json:9-16``):
    linear from ``warmup_min_ratio`` to 1 over the warm-up steps, then half a cosine down to ``cos_min_ratio`` at
    ``total_num_steps``, constant afterwards (unlike ``CosineAnnealingLR`` it is not periodic)."""
    if step < warmup_num_steps:
        return warmup_min_ratio + (1.0 - warmup_min_ratio) * step / max(warmup_num_steps, 1)
    span = max(total_num_steps - warmup_num_steps, 1)
    t = min(step - warmup_num_steps, span) / span
    return cos_min_ratio + (1.0 - cos_min_ratio) * (1 + math.cos(math.pi * t)) / 2


def make_warmup_cosine(optimizer: torch.optim.Optimizer, **params):
    """``LambdaLR`` over :func:`warmup_cosine_ratio`; ``params`` are the json's ``scheduler.params``."""
    keys = ("total_num_steps", "warmup_num_steps", "cos_min_ratio", "warmup_min_ratio")
    kw = {k: params[k] for k in keys if k in params}
    return torch.optim.lr_scheduler.LambdaLR(optimizer, lambda step: warmup_cosine_ratio(step, **kw))


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
