"""Metrics record, sinks and logging setup.

The record carries the reference's keys (``01-single-gpu/train_llm.py:156-176``):
``global_step, lr, running_loss, epoch, epoch_progress, num_batches_remaining, total_gb, curr_alloc_gb,
peak_alloc_gb, curr_resv_gb, peak_resv_gb, tok/s, time/total, time/{data,forward,backward,update}``.
It is logged on every rank and sent to wandb from rank 0 (``02-.../train_llm.py:171-197``). The box is
offline, so wandb is optional and a JSONL file with the same keys is always written; the alternative
wandb grouping modes of ``related-topics/wandb-configurations/README.md:22-63`` are ``--wandb-mode``.
"""
from __future__ import annotations

import json
import logging
import os
from pathlib import Path
from typing import Dict, Optional

from .timers import get_mem_stats

LOGGER = logging.getLogger("b200train")


def setup_logging(rank: int = 0, level=logging.INFO):
    """``[rank=R] [time] LEVEL:msg`` (reference: ``02-distributed-data-parallel/train_llm.py:43-46``)."""
    logging.basicConfig(format=f"[rank={rank}] [%(asctime)s] %(levelname)s:%(message)s", level=level, force=True)
    return LOGGER


def build_record(state: dict, lr: float, log_freq: int, num_batches: int, i_step: int, tok_per_step: int,
                 timers: dict, device) -> Dict[str, float]:
    ms_per_step = sum(t.avg_elapsed_ms() for k, t in timers.items() if k != "waiting")
    info = {
        "global_step": state["global_step"],
        "lr": lr,
        "running_loss": state["running_loss"] / log_freq,
        "epoch": state["epoch"],
        "epoch_progress": state["epoch_step"] / max(num_batches, 1),
        "num_batches_remaining": num_batches - i_step,
        **get_mem_stats(device),
        "tok/s": 1000 * tok_per_step / ms_per_step if ms_per_step > 0 else 0.0,
        "time/total": ms_per_step,
        **{f"time/{k}": t.avg_elapsed_ms() for k, t in timers.items()},
    }
    return info


class MetricsSink:
    """JSONL file (always) + wandb (when importable and not disabled)."""

    def __init__(self, exp_dir: Path, rank: int, local_rank: int = 0, experiment_name: str = "",
                 resumed: bool = False, config: Optional[dict] = None, mode: str = "rank0"):
        self.rank = rank
        self.run = None
        self.path = Path(exp_dir) / ("metrics.jsonl" if rank == 0 else f"metrics-rank{rank}.jsonl")
        want_wandb = {"rank0": rank == 0, "local_rank0": local_rank == 0, "all": True, "off": False}[mode]
        self.write_file = rank == 0 or mode in ("all",) or (mode == "local_rank0" and local_rank == 0)
        if want_wandb and os.environ.get("WANDB_MODE", "").lower() != "disabled":
            try:
                import wandb

                kw = dict(project="distributed-training-guide", dir=str(exp_dir), name=experiment_name,
                          id=experiment_name, resume="must" if resumed else None, save_code=True, config=config)
                if mode != "rank0":
                    kw.update(group=experiment_name, name=f"{experiment_name}-rank{rank}",
                              id=f"{experiment_name}-rank{rank}")
                self.run = wandb.init(**kw)
            except Exception as e:  # offline box / no login: metrics still go to the JSONL file
                LOGGER.warning("wandb unavailable (%s); metrics are written to %s", e, self.path)

    def log(self, info: dict, step: int):
        if self.write_file:
            with open(self.path, "a") as fp:
                fp.write(json.dumps(info) + "\n")
        if self.run is not None:
            self.run.log(info, step=step)

    def close(self):
        if self.run is not None:
            self.run.finish()
