"""DataLoader + sampler construction.

ch01: ``DataLoader(shuffle=True, drop_last=True, collate_fn=default_data_collator)``
(``01-single-gpu/train_llm.py:56-62``); ch02/04: ``DistributedSampler(shuffle=True, drop_last=True)`` +
``set_epoch`` (``02-.../train_llm.py:72-78,143``); ch05: ``num_workers=1, prefetch_factor=2``
(``05-.../train_llm.py:134-142``); ch06/07: the sampler spans the **dp** sub-mesh only, so all ranks of a
tp group read the same batch (``06-tensor-parallel/train_llm.py:152-158``). Unlike ch06/07
``set_epoch`` is called for every strategy.
"""
from __future__ import annotations

import random

import numpy as np
import torch
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler


def default_data_collator(features):
    """Stack a list of ``{name: tensor | list}`` rows into ``{name: [B, ...] tensor}``."""
    out = {}
    for k in features[0].keys():
        vals = [f[k] if torch.is_tensor(f[k]) else torch.as_tensor(f[k]) for f in features]
        out[k] = torch.stack(vals)
    return out


def seed_worker(worker_id):
    """Determinism recipe of ``related-topics/determinism/README.md:30-40``."""
    s = torch.initial_seed() % 2 ** 32
    np.random.seed(s)
    random.seed(s)


def make_dataloader(dataset, batch_size: int, dp_rank: int = 0, dp_size: int = 1, seed: int = 0,
                    num_workers: int = 0, prefetch_factor=None, collate_fn=None, shuffle: bool = True,
                    pin_memory: bool = False, distributed: bool = None):
    distributed = (dp_size > 1) if distributed is None else distributed
    gen = torch.Generator().manual_seed(seed)
    kw = dict(batch_size=batch_size, collate_fn=collate_fn or default_data_collator, num_workers=num_workers,
              pin_memory=pin_memory, worker_init_fn=seed_worker, generator=gen)
    if num_workers > 0 and prefetch_factor:
        kw["prefetch_factor"] = prefetch_factor
    if distributed:
        sampler = DistributedSampler(dataset, num_replicas=dp_size, rank=dp_rank, shuffle=shuffle, seed=seed,
                                     drop_last=True)
        return DataLoader(dataset, sampler=sampler, drop_last=True, **kw)
    return DataLoader(dataset, shuffle=shuffle, drop_last=True, **kw)


def set_epoch(dataloader, epoch: int):
    sampler = getattr(dataloader, "sampler", None)
    if hasattr(sampler, "set_epoch"):
        sampler.set_epoch(epoch)
