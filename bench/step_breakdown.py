"""Per-kernel device time of one training step (torch.profiler / CUPTI activity records; diagnostic only,
the timed numbers come from bench.py). Prints the busy time per kernel name, the span of the step on the
device and the idle time between kernels on the compute stream."""
import argparse
import collections
import json
import sys
from pathlib import Path

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200train.models import load_config  # noqa: E402
from b200train.parallel import make_strategy  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--seq", type=int, default=4096)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--fuse", default="stream")
ap.add_argument("--out", default="gpurun_out/step_breakdown.json")
args = ap.parse_args()
cfg = load_config(args.model)
dev = torch.device("cuda:0")
eng = make_strategy("fsdp", cfg, dev, lr=3e-5, seed=0, fuse_optimizer=False if args.fuse == "off" else args.fuse)
ids = torch.randint(0, cfg.vocab_size, (args.batch, args.seq), device=dev)
batch = {"input_ids": ids, "labels": ids}
for _ in range(3):
    eng.train_step(batch)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    eng.train_step(batch)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
by = collections.defaultdict(lambda: [0, 0.0])
ivals = []
for e in evs:
    name = e.name.split("<")[0].split("(")[0][:60]
    by[name][0] += 1
    by[name][1] += e.device_time
    ivals.append((e.time_range.start, e.time_range.end, name))
ivals.sort()
span = (max(b for _, b, _ in ivals) - ivals[0][0]) / 1e3
busy, cur_s, cur_e, cur_n = 0.0, *ivals[0]
gaps, gap_where = [], []
for s, e_, n_ in ivals[1:]:
    if s > cur_e:
        busy += cur_e - cur_s
        gaps.append(s - cur_e)
        gap_where.append((s - cur_e, cur_n, n_))
        cur_s, cur_e, cur_n = s, e_, n_
    else:
        if e_ > cur_e:
            cur_e, cur_n = e_, n_
busy += cur_e - cur_s
rows = sorted(((v[1] / 1e3, v[0], k) for k, v in by.items()), reverse=True)
out = {"model": args.model, "span_ms": span, "union_busy_ms": busy / 1e3, "idle_ms": span - busy / 1e3,
       "n_kernels": len(evs), "n_gaps": len(gaps), "median_gap_us": sorted(gaps)[len(gaps) // 2] if gaps else 0,
       "kernels": [{"ms": round(ms, 3), "n": n, "name": k} for ms, n, k in rows[:40]]}
Path(args.out).parent.mkdir(exist_ok=True)
Path(args.out).write_text(json.dumps(out, indent=1))
print(json.dumps({k: v for k, v in out.items() if k != "kernels"}))
for ms, n, k in rows[:25]:
    print(f"{ms:9.3f} ms  {n:5d}  {k}")
print("largest idle gaps (us, kernel before -> kernel after):")
for g, a, b in sorted(gap_where, reverse=True)[:12]:
    print(f"{g:9.1f}  {a} -> {b}")
