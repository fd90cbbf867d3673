"""Per-experiment resume state and the single-file checkpoint flavours.

Layout under ``exp_dir = <save-dir>/<experiment-name>`` (reference: ``01-single-gpu/train_llm.py:72-93,183-189``):
  ``state.json``       {"epoch", "global_step", "epoch_step", "running_loss"} -- its presence triggers resume
  ``lr_scheduler.pt``  torch.save(lr_scheduler.state_dict())
  ``model.pt``, ``optimizer.pt``   full state dicts (chapter 01; chapter 02 saves the model only,
                                   ``02-distributed-data-parallel/train_llm.py:199-206``)
  ``rng.pt``           numpy / random / torch / cuda RNG states (``related-topics/determinism/README.md:46-68``)
``state.json`` is written last and atomically (temp file + rename), which closes the reference's
"crash between shards and state.json" window for the single-file flavours.
"""
from __future__ import annotations

import json
import os
import random
from pathlib import Path

import numpy as np
import torch


def new_state() -> dict:
    return {"epoch": 0, "global_step": 0, "epoch_step": 0, "running_loss": 0}


def has_checkpoint(exp_dir: Path) -> bool:
    return (Path(exp_dir) / "state.json").exists()


def load_state(exp_dir: Path) -> dict:
    with open(Path(exp_dir) / "state.json") as fp:
        return json.load(fp)


def save_state(exp_dir: Path, state: dict):
    exp_dir = Path(exp_dir)
    tmp = exp_dir / "state.json.tmp"
    with open(tmp, "w") as fp:
        json.dump(state, fp)
    os.replace(tmp, exp_dir / "state.json")


def save_full(exp_dir: Path, model_sd=None, optimizer_sd=None, lr_scheduler=None, state=None):
    """Flavour A / B: plain ``torch.save`` files, then ``state.json``."""
    exp_dir = Path(exp_dir)
    exp_dir.mkdir(parents=True, exist_ok=True)
    if optimizer_sd is not None:
        _save_atomic(optimizer_sd, exp_dir / "optimizer.pt")
    if model_sd is not None:
        _save_atomic(model_sd, exp_dir / "model.pt")
    if lr_scheduler is not None:
        _save_atomic(lr_scheduler.state_dict(), exp_dir / "lr_scheduler.pt")
    if state is not None:
        save_state(exp_dir, state)


def _save_atomic(obj, path: Path):
    """``torch.save`` to a temporary name, then rename: a crash in the middle of a checkpoint leaves the previous
    file whole (``state.json`` still describes it) instead of a torn one."""
    tmp = path.with_name(path.name + ".tmp")
    torch.save(obj, tmp)
    os.replace(tmp, path)


def load_full(exp_dir: Path, device, want_optimizer: bool = True):
    exp_dir = Path(exp_dir)
    out = {"model": torch.load(exp_dir / "model.pt", map_location=device, weights_only=True)}
    opt = exp_dir / "optimizer.pt"
    if want_optimizer and opt.exists():  # chapter 02 of the reference never writes it
        out["optimizer"] = torch.load(opt, map_location=device, weights_only=False)
    lr = exp_dir / "lr_scheduler.pt"
    if lr.exists():
        out["lr_scheduler"] = torch.load(lr, map_location=device, weights_only=True)
    out["state"] = load_state(exp_dir)
    return out


def save_rng(exp_dir: Path, rank: int = 0):
    torch.save({
        "np": np.random.get_state(),
        "random": random.getstate(),
        "torch": torch.get_rng_state(),
        "cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else [],
    }, Path(exp_dir) / ("rng.pt" if rank == 0 else f"rng-rank{rank}.pt"))


def load_rng(exp_dir: Path, rank: int = 0):
    p = Path(exp_dir) / ("rng.pt" if rank == 0 else f"rng-rank{rank}.pt")
    if not p.exists():
        return False
    s = torch.load(p, weights_only=False)
    np.random.set_state(s["np"])
    random.setstate(s["random"])
    torch.set_rng_state(s["torch"])
    if torch.cuda.is_available() and s["cuda"]:
        torch.cuda.set_rng_state_all(s["cuda"])
    return True


def set_determinism(seed: int):
    """``related-topics/determinism/README.md:16-28``."""
    torch.manual_seed(seed)
    np.random.seed(seed)
    random.seed(seed)
    torch.use_deterministic_algorithms(True, warn_only=True)
    os.environ.setdefault("CUBLAS_WORKSPACE_CONFIG", ":4096:8")
