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

    # ---- checkpoint surface (ckpt/dcp.py walks units through these) --------------------------------
    def num_units(self) -> int:
        return len(self.units)

    def optimizer_param_groups(self):
        g = {k: v for k, v in self.optimizer.param_groups[0].items() if k != "params"}
        g["betas"] = list(g["betas"])
        return [g]

    def set_optimizer_param_groups(self, groups):
        """Write checkpointed hyper-parameters (lr, betas, eps, weight_decay, initial_lr) back into the live
        optimizer: torch LR schedulers continue recursively from ``param_groups[0]['lr']``, so a resume that
        leaves the freshly constructed lr in place walks a different schedule."""
        if not groups:
            return
        src = groups[0]
        dst = self.optimizer.param_groups[0]
        for k in ("lr", "betas", "eps", "weight_decay", "initial_lr"):
            if k in src:
                v = src[k]
                if isinstance(v, torch.Tensor):
                    v = v.item() if v.numel() == 1 else v.tolist()
                dst[k] = tuple(v) if k == "betas" else float(v)

    def optimizer_step_count(self) -> int:
        return self.optimizer._step

    def set_optimizer_step_count(self, n: int):
        self.optimizer._step = n

    def _unit_full_flat(self, i: int, kind: str) -> torch.Tensor:
        """Full (unsharded) flat tensor of unit ``i``: parameters or one AdamW moment."""
        if kind == "param":
            return self.units[i].flat_param
        return (self.optimizer.exp_avg if kind == "exp_avg" else self.optimizer.exp_avg_sq)[i]

    def _unit_store_flat(self, i: int, kind: str, full: torch.Tensor):
        self._unit_full_flat(i, kind).copy_(full)

    def _named_views(self, i: int, flat: torch.Tensor):
        """HF-named views of a flat tensor laid out like unit ``i``."""
        u = self.units[i]
        saved = [p.data for p in u.params]
        for p, o in zip(u.params, u.offsets):
            p.data = flat[o:o + p.numel()].view(p.shape)
        lo = flat.data_ptr()
        hi = lo + flat.numel() * flat.element_size()
        out = {}
        for k, v in self.model.hf_state_dict().items():
            if lo <= v.data_ptr() < hi and not (k == "lm_head.weight" and self.cfg.tie_word_embeddings):
                out[k] = v
        for p, d in zip(u.params, saved):
            p.data = d
        return out

    @torch.no_grad()
    def unit_named_full(self, i: int, kind: str, names_only: bool = False):
        if names_only:
            return self._named_views(i, self._names_probe(i)).keys()
        return self._named_views(i, self._unit_full_flat(i, kind))

    def _names_probe(self, i: int):
        """Scratch laid out like unit ``i`` (only its addresses are used: which named tensor sits where). Not kept:
        a unit of Llama-405B is 6.4 GB, and checkpoints are rare."""
        return torch.empty(self.units[i].numel, dtype=self.dtype, device=self.device)

    @torch.no_grad()
    def unit_store_named(self, i: int, kind: str, tensors):
        u = self.units[i]
        flat = torch.zeros(u.numel, dtype=self.dtype, device=self.device)
        for name, view in self._named_views(i, flat).items():
            view.copy_(tensors[name].to(device=self.device, dtype=self.dtype))
        self._unit_store_flat(i, kind, flat)
