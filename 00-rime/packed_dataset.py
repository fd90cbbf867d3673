#!/usr/bin/env python
"""Chapter 00 (fork): packed-sequence collator and the data-loading throughput benchmark.

Importing this module gives the collator under the names the reference's chapter uses
(``00-rime/packed_dataset.py:11-49``); the implementation lives in ``b200train/data/packed.py``.
Running it is the benchmark of ``00-rime/packed_dataset.py:51-88``::

    python packed_dataset.py                                        # one process
    torchrun --standalone --nproc-per-node 8 packed_dataset.py      # one loader per rank

Each rank prints its first batch (the ranks must not see the same rows) and rank 0 the global tokens per
second every 50 iterations.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")))

from b200train.data.packed import (  # noqa: E402,F401
    block_diagonal_causal_mask,
    cu_seqlens_from_positions,
    packed_positions,
    pad_and_collate,
)
from b200train.data.packed_bench import run  # noqa: E402

# the reference's name for the position-id / mask helper
get_attention_mask_and_pos_ids_for_packed_sequence = packed_positions


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seq-length", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=1000)
    ap.add_argument("--num-workers", type=int, default=8)
    a = ap.parse_args()
    run(a.seq_length, a.iters, a.num_workers)
