"""How far ahead of the GPU does the host run? Times the host-side enqueue of one training step (no sync
inside) against the device time of the same step. enqueue << device: the step is GPU-bound and launch
gaps are hidden; enqueue ~ device: the GPU waits for launches."""
import argparse
import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200train.models import load_config  # noqa: E402
from b200train.parallel import make_strategy  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--seq", type=int, default=4096)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--fuse", default="stream")
args = ap.parse_args()
cfg = load_config(args.model)
dev = torch.device("cuda:0")
eng = make_strategy("fsdp", cfg, dev, lr=3e-5, seed=0, fuse_optimizer=False if args.fuse == "off" else args.fuse)
ids = torch.randint(0, cfg.vocab_size, (args.batch, args.seq), device=dev)
batch = {"input_ids": ids, "labels": ids}
for _ in range(3):
    eng.train_step(batch)
torch.cuda.synchronize()
enq, devms = [], []
for _ in range(args.steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    eng.train_step(batch)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    enq.append((t1 - t0) * 1e3)
    devms.append(e0.elapsed_time(e1))
print(json.dumps({"model": args.model, "seq": args.seq, "batch": args.batch, "host_enqueue_ms": sorted(enq)[len(enq) // 2],
                  "device_ms": sorted(devms)[len(devms) // 2]}))
