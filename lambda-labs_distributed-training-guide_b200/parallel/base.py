"""Strategy interface shared by the single-device, DDP, FSDP, TP and 2D engines.

The reference wraps an HF module in a PyTorch parallel wrapper and then runs the same loop
(``outputs = model(**batch)``, ``loss.backward()``, ``optimizer.step()``, ``lr_scheduler.step()``:
``01-single-gpu/train_llm.py:139-149``). Here a strategy owns the model's flat storage, the comm
schedule and the optimizer, and exposes those same four phases.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from ..models import ModelConfig, build_model
from ..utils.flat import FlatUnit
from ..utils.schedule import make_lr_scheduler
from .optimizer import FlatAdamW


class Strategy:
    name = "base"

    def __init__(self, cfg: ModelConfig, device, lr: float = 3e-5, seed: int = 0, dtype=torch.bfloat16):
        self.cfg = cfg
        self.device = torch.device(device)
        self.lr = lr
        self.seed = seed
        self.dtype = dtype
        self.model = None
        self.units = []
        self.optimizer: Optional[FlatAdamW] = None
        self.lr_scheduler = None

    # ---- construction helpers -------------------------------------------------------------------
    def _build_local_model(self):
        torch.manual_seed(self.seed)
        model = build_model(self.cfg, dtype=self.dtype, device=self.device, seed=self.seed)
        return model

    def _flatten(self, model, pad_to: int = 1):
        return [FlatUnit(name, params, pad_to=pad_to) for name, params in model.units()]

    def _make_optimizer(self, **kw):
        self.optimizer = FlatAdamW(self.units, lr=self.lr, **kw)
        self.lr_scheduler = make_lr_scheduler(self.optimizer, self.lr)

    # ---- the four phases ------------------------------------------------------------------------
    def to_device(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        return {k: v.to(device=self.device, non_blocking=True) for k, v in batch.items()}

    def forward(self, batch):
        return self.model(**batch)

    def backward(self, loss: torch.Tensor):
        loss.backward()

    def zero_grad(self):
        self.optimizer.zero_grad()

    def optimizer_step(self):
        self.optimizer.step()
        self.lr_scheduler.step()

    def train_step(self, batch) -> torch.Tensor:
        """H2D copy + forward + backward + update; returns the (device) loss of this step."""
        batch = self.to_device(batch)
        out = self.forward(batch)
        self.zero_grad()
        self.backward(out.loss)
        self.optimizer_step()
        return out.loss.detach()

    # ---- introspection ---------------------------------------------------------------------------
    def num_parameters(self) -> int:
        return self.cfg.num_parameters()

    def tokens_per_step(self, batch_size: int, seq_len: int) -> int:
        """Global tokens per optimizer step (reference: ``world_size * batch * seq``, ``02-.../train_llm.py:172``)."""
        return batch_size * seq_len

    def barrier(self):
        pass

    @property
    def rank(self):
        return 0

    @property
    def world_size(self):
        return 1

    def model_state_dict(self):
        return self.model.hf_state_dict()

    def load_model_state_dict(self, sd):
        self.model.load_hf_state_dict(sd)
