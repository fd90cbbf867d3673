"""Per-stream timeline of one FSDP training step on N GPUs (rank 0's view; diagnostic only).

Launch with ``python -m torch.distributed.run --nproc-per-node N bench/step_timeline.py`` (or plain
``python`` for one GPU). One step is recorded with torch.profiler (CUPTI activity records), exported as
a chrome trace and folded per CUDA stream: busy time, idle gaps on the compute stream with the kernels
on either side, per-kernel totals, and what the communication stream was doing. Comparing the N=1 and
N>1 outputs shows where the multi-GPU step loses time: slower GEMMs (contention), exposed collectives
(gaps), or the unhidden head / tail of the step.
"""
import argparse
import collections
import json
import os
import sys
import tempfile
from pathlib import Path

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200train.models import load_config  # noqa: E402
from b200train.parallel import make_strategy  # noqa: E402
from b200train.parallel.comm import init_distributed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--seq", type=int, default=4096)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--layers", type=int, default=None)
ap.add_argument("--fuse", default="stream")
ap.add_argument("--out", default="gpurun_out/step_timeline.json")
ap.add_argument("--cpu", action="store_true", help="also record CUDA runtime calls (host side) around the gaps")
ap.add_argument("--windows", type=int, default=4, help="dump every event around this many of the largest gaps")
args = ap.parse_args()
rank, local_rank, world, dev = init_distributed("cuda")
cfg = load_config(args.model)
if args.layers:
    cfg.num_hidden_layers = args.layers
eng = make_strategy("fsdp", cfg, dev, lr=3e-5, seed=0, fuse_optimizer=False if args.fuse == "off" else args.fuse)
ids = torch.randint(0, cfg.vocab_size, (args.batch, args.seq), device=dev)
batch = {"input_ids": ids, "labels": ids}
for _ in range(3):
    eng.train_step(batch)
torch.cuda.synchronize()
if world > 1:
    torch.distributed.barrier()
acts = [ProfilerActivity.CUDA] + ([ProfilerActivity.CPU] if args.cpu else [])
with profile(activities=acts) as prof:
    eng.train_step(batch)
    eng.train_step(batch)
    torch.cuda.synchronize()
tmp = tempfile.NamedTemporaryFile(suffix=".json", delete=False).name
prof.export_chrome_trace(tmp)
trace = json.load(open(tmp))
os.unlink(tmp)
evs = []
host = []
for e in trace["traceEvents"]:
    if e.get("ph") == "X" and e.get("cat") in ("cuda_runtime", "cuda_driver"):
        host.append((float(e["ts"]), float(e["ts"]) + float(e["dur"]), e["name"][:40]))
    if e.get("ph") != "X" or e.get("cat") not in ("kernel", "gpu_memcpy", "gpu_memset"):
        continue
    evs.append((float(e["ts"]), float(e["ts"]) + float(e["dur"]), e["name"].split("<")[0].split("(")[0][:56],
                int(e.get("args", {}).get("stream", -1))))
evs.sort()
# the second recorded step: from the first kernel after the midpoint gap... simpler: split at the embedding forward
emb = [i for i, e in enumerate(evs) if "embedding_fwd" in e[2]]
start_i = emb[1] if len(emb) > 1 else 0
# include the collectives issued just before the second step's embedding (barrier + first gathers)
t_first = evs[emb[0]][0]
step = [e for e in evs if e[0] >= evs[start_i][0] - 0] if len(emb) > 1 else evs
prev = [e for e in evs if e[0] < evs[start_i][0]]
step_span = (step[-1][1] if step else 0) - (step[0][0] if step else 0)
period = evs[emb[1]][0] - evs[emb[0]][0] if len(emb) > 1 else step_span
streams = collections.Counter(e[3] for e in step)
main = streams.most_common(1)[0][0]
out = {"world": world, "period_ms": period / 1e3, "streams": {}}
for s, _ in streams.most_common():
    iv = [(a, b, n) for a, b, n, st in step if st == s]
    busy = 0.0
    cur_a, cur_b = iv[0][0], iv[0][1]
    for a, b, _ in iv[1:]:
        if a > cur_b:
            busy += cur_b - cur_a
            cur_a, cur_b = a, b
        else:
            cur_b = max(cur_b, b)
    busy += cur_b - cur_a
    by = collections.defaultdict(lambda: [0, 0.0])
    for a, b, n in iv:
        by[n][0] += 1
        by[n][1] += b - a
    out["streams"][str(s)] = {"kernels": len(iv), "busy_ms": busy / 1e3,
                              "top": [{"ms": round(v[1] / 1e3, 3), "n": v[0], "name": k}
                                      for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])[:14]]}
iv = [(a, b, n) for a, b, n, st in step if st == main]
gaps = []
for (a0, b0, n0), (a1, b1, n1) in zip(iv, iv[1:]):
    if a1 > b0:
        gaps.append((a1 - b0, n0, n1, (a1 - iv[0][0]) / 1e3))
out["main_stream"] = main
out["main_idle_ms"] = sum(g[0] for g in gaps) / 1e3
out["gaps_over_50us_ms"] = sum(g[0] for g in gaps if g[0] > 50) / 1e3
out["largest_gaps"] = [{"us": round(g[0], 1), "after": g[1], "before": g[2], "at_ms": round(g[3], 2)}
                       for g in sorted(gaps, reverse=True)[:16]]
# the tail: last compute-stream kernel of the first recorded step -> first kernel of the second step
if len(emb) > 1:
    last_main_prev = max((e[1] for e in prev if e[3] == main), default=None)
    last_any_prev = max((e[1] for e in prev), default=None)
    out["tail"] = {"main_stream_idle_before_step_us": round(iv[0][0] - last_main_prev, 1) if last_main_prev else None,
                   "all_streams_idle_before_step_us": round(iv[0][0] - last_any_prev, 1) if last_any_prev else None}
if rank == 0:
    Path(args.out).parent.mkdir(exist_ok=True)
    Path(args.out).write_text(json.dumps(out, indent=1))
    print(json.dumps({k: v for k, v in out.items() if k not in ("streams", "largest_gaps")}))
    for s, d in out["streams"].items():
        print(f"--- stream {s}: {d['kernels']} kernels, busy {d['busy_ms']:.2f} ms")
        for r in d["top"]:
            print(f"{r['ms']:9.3f} ms {r['n']:5d}  {r['name']}")
    print("largest gaps on the compute stream:")
    for g in out["largest_gaps"]:
        print(f"{g['us']:9.1f} us at {g['at_ms']:7.2f} ms  {g['after']} -> {g['before']}")
if rank == 0 and args.windows:
    # everything that ran on any stream (and, with --cpu, every runtime call) around the largest gaps
    big = sorted(gaps, reverse=True)
    picked = []
    for g in big:
        key = (g[1], g[2])
        if sum(1 for k in picked if k == key) < 2:
            picked.append(key)
            t_gap_end = iv[0][0] + g[3] * 1e3
            lo, hi = t_gap_end - g[0] - 400, t_gap_end + 150
            print(f"=== window around gap of {g[0]:.1f} us at {g[3]:.2f} ms ({g[1]} -> {g[2]}); t=0 is the gap end")
            for a, b, n, st in step:
                if b >= lo and a <= hi:
                    print(f"   gpu s{st:<3d} {a - t_gap_end:9.1f} .. {b - t_gap_end:9.1f}  {n}")
            for a, b, n in host:
                if b >= lo and a <= hi and b - a > 15:
                    print(f"   host      {a - t_gap_end:9.1f} .. {b - t_gap_end:9.1f}  {n}")
        if len(picked) >= args.windows:
            break
if world > 1:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
