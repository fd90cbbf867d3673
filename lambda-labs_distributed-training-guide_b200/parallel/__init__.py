"""Parallel strategies: single, ddp (+zero1), fsdp (+offload, +activation
checkpointing), tp (+sequence / loss parallel) and 2d."""
from .base import Strategy
from .single import SingleDevice


def make_strategy(name: str, cfg, device, **kw) -> Strategy:
    name = name.lower()
    if name == "single":
        return SingleDevice(cfg, device, **kw)
    if name == "ddp":
        from .ddp import DDP
        return DDP(cfg, device, **kw)
    if name == "fsdp":
        from .fsdp import FSDP
        return FSDP(cfg, device, **kw)
    if name in ("hsdp", "hybrid_shard"):
        from .fsdp import HSDP
        return HSDP(cfg, device, **kw)
    if name == "tp":
        from .tp import TensorParallel
        return TensorParallel(cfg, device, **kw)
    if name in ("2d", "fsdp_tp"):
        from .tp import TwoD
        return TwoD(cfg, device, **kw)
    raise ValueError(f"unknown strategy {name}")


__all__ = ["Strategy", "SingleDevice", "make_strategy"]
