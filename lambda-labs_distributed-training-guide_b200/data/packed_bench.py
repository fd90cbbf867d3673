"""Data-loading throughput benchmark (``00-rime/packed_dataset.py:51-88`` ``__main__``): optional process
group, ``DataLoader(batch_size=1, num_workers=8, DistributedSampler(shuffle=False))``, 1000 iterations,
first batch printed per rank (no duplication across ranks), global tok/s every 50 iterations.
Published reference numbers: ~90k tok/s for 1 process, ~600k for 8 (``00-rime/README.md:7-27``)."""
from __future__ import annotations

import argparse
import os
import time


from .datasets import SyntheticTokens
from .loader import make_dataloader
from .packed import pad_and_collate


def run(seq_length=8192, iters=1000, num_workers=8, vocab=128_263, eos_every=701, log_every=50, quiet=False):
    import torch.distributed as dist

    world, rank = 1, 0
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        if not dist.is_initialized():
            dist.init_process_group("gloo")
        world, rank = dist.get_world_size(), dist.get_rank()

    class Packed(SyntheticTokens):
        def __getitem__(self, i):
            row = super().__getitem__(i)
            ids = row["input_ids"]
            ids[eos_every - 1::eos_every] = 128_262
            return {"input_ids": ids}

    ds = Packed(vocab - 1, seq_length, n_samples=max(iters * world, world))
    dl = make_dataloader(ds, 1, dp_rank=rank, dp_size=world, num_workers=num_workers, shuffle=False,
                         collate_fn=pad_and_collate, distributed=True)
    t0 = time.time()
    rates = []
    for i, batch in enumerate(dl):
        if i == 0 and not quiet:
            print(f"[rank={rank}] first batch {batch['input_ids'][0, :8].tolist()}")
        if (i + 1) % log_every == 0:
            rate = world * (i + 1) * seq_length / (time.time() - t0)
            rates.append(rate)
            if rank == 0 and not quiet:
                print(f"iter {i + 1}: {rate:,.0f} tok/s")
        if i + 1 >= iters:
            break
    return rates


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seq-length", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=1000)
    ap.add_argument("--num-workers", type=int, default=8)
    a = ap.parse_args()
    run(a.seq_length, a.iters, a.num_workers)
