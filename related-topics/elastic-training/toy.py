#!/usr/bin/env python
"""CPU-only fault-injection toy for elastic restarts (no GPU required).

Every rank draws a random number per step and raises when it falls below ``--fail-prob`` (the reference
uses 0.001, ``related-topics/elastic-training/toy.py:30-36``); rank 0 persists the step counter between
barriers so a restarted gang resumes where the last one stopped. Run under
``torchrun --nnodes=1 --nproc-per-node 4 --max-restarts 3 toy.py`` and read the per-attempt ``error.json``.
``--fail-at-step N`` / ``--fail-rank R`` make the failure deterministic (used by the tests).
"""
import argparse
import json
import os
import random
from pathlib import Path

import torch.distributed as dist

try:
    from torch.distributed.elastic.multiprocessing.errors import record
except Exception:  # pragma: no cover
    def record(fn):
        return fn


@record
def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--state-file", default="./toy-state.json")
    ap.add_argument("--total-steps", type=int, default=1000)
    ap.add_argument("--fail-prob", type=float, default=0.001)
    ap.add_argument("--fail-at-step", type=int, default=None)
    ap.add_argument("--fail-rank", type=int, default=0)
    args = ap.parse_args(argv)
    import datetime

    dist.init_process_group("gloo", timeout=datetime.timedelta(seconds=60))
    rank, world = dist.get_rank(), dist.get_world_size()
    path = Path(args.state_file)
    state = json.loads(path.read_text()) if path.exists() else {"num_steps": 0}
    restart = int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"))
    print(f"[rank={rank}] world={world} restart={restart} resuming at step {state['num_steps']}", flush=True)
    while state["num_steps"] < args.total_steps:
        random.seed(rank + world * state["num_steps"])
        value = random.random()
        deterministic = (args.fail_at_step is not None and state["num_steps"] == args.fail_at_step
                         and rank == args.fail_rank and restart == 0)
        if deterministic or (args.fail_at_step is None and value < args.fail_prob):
            raise ValueError(f"injected failure on rank {rank} at step {state['num_steps']}")
        state["num_steps"] += 1
        dist.barrier()
        if rank == 0:
            path.write_text(json.dumps(state))
        dist.barrier()
    print(f"[rank={rank}] finished {state['num_steps']} steps", flush=True)
    dist.destroy_process_group()
    return state


if __name__ == "__main__":
    main()
