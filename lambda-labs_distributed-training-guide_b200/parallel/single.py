"""Single-device strategy (chapter 01: ``01-single-gpu/train_llm.py``; runs on CPU too)."""
from __future__ import annotations

from .base import Strategy


class SingleDevice(Strategy):
    name = "single"

    def __init__(self, cfg, device, lr=3e-5, seed=0, **kw):
        super().__init__(cfg, device, lr, seed, **kw)
        self.model = self._build_local_model()
        self.units = self._flatten(self.model)
        self._make_optimizer()
