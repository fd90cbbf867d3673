#!/usr/bin/env python
"""Cluster GPU monitor: polls every host in a hosts file over ssh with ``nvidia-smi`` and prints per-node
and cluster-wide averages of utilisation, power (% of limit), memory (% used) and the number of compute
processes -- a falling process count or power near 10 % flags a hung job.

Same command line as the reference tool (``top-cluster.py hosts [--poll-freq 1000]``,
``top-cluster.py:6-11``); the parsing is factored out (``parse_node_output``) so it is unit-tested on
canned ``nvidia-smi`` text. ``--once`` prints a single table; ``--local`` queries this machine without ssh.
"""
import argparse
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor

QUERY = ("nvidia-smi --query-gpu=utilization.gpu,power.draw,power.limit,memory.used,memory.total "
         "--format=csv,noheader,nounits && nvidia-smi --query-compute-apps=pid --format=csv,noheader")


def parse_node_output(text: str) -> dict:
    """GPU rows contain commas (5 numeric fields); process rows are bare pids."""
    gpus, procs = [], 0
    for line in text.strip().splitlines():
        line = line.strip()
        if not line:
            continue
        if "," in line:
            try:
                util, draw, limit, used, total = [float(x) for x in line.split(",")]
            except ValueError:
                continue
            gpus.append({"util": util, "power": 100.0 * draw / limit if limit else 0.0,
                         "memory": 100.0 * used / total if total else 0.0})
        else:
            procs += 1
    n = max(len(gpus), 1)
    return {"ngpus": len(gpus), "nprocs": procs,
            "util": sum(g["util"] for g in gpus) / n,
            "power": sum(g["power"] for g in gpus) / n,
            "memory": sum(g["memory"] for g in gpus) / n}


def query_host(host: str, local: bool = False, timeout: int = 20) -> dict:
    cmd = ["bash", "-c", QUERY] if local else ["ssh", "-o", "BatchMode=yes", "-o", "ConnectTimeout=5", host, QUERY]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
        if out.returncode != 0:
            return {"error": (out.stderr or "failed").strip().splitlines()[-1][:60]}
        return parse_node_output(out.stdout)
    except Exception as e:  # unreachable host, timeout
        return {"error": str(e)[:60]}


def summarize(stats: dict) -> dict:
    ok = [s for s in stats.values() if "error" not in s and s["ngpus"]]
    n = max(len(ok), 1)
    return {"util": sum(s["util"] for s in ok) / n, "power": sum(s["power"] for s in ok) / n,
            "memory": sum(s["memory"] for s in ok) / n, "nprocs": sum(s["nprocs"] for s in ok),
            "nodes": len(ok)}


def render(stats: dict) -> str:
    lines = [f"{'name':<24}{'util':>8}{'power':>8}{'memory':>8}{'nprocs':>8}", "-" * 56]
    c = summarize(stats)
    lines.append(f"{'cluster':<24}{c['util']:>7.1f}%{c['power']:>7.1f}%{c['memory']:>7.1f}%{c['nprocs']:>8}")
    lines.append("-" * 56)
    for host, s in stats.items():
        if "error" in s:
            lines.append(f"{host:<24}  unreachable: {s['error']}")
        else:
            lines.append(f"{host:<24}{s['util']:>7.1f}%{s['power']:>7.1f}%{s['memory']:>7.1f}%{s['nprocs']:>8}")
    return "\n".join(lines)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("hosts", nargs="?", default=None, help="file with one hostname per line")
    ap.add_argument("--poll-freq", type=int, default=1000, help="milliseconds between polls")
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--local", action="store_true", help="query this machine only (no ssh)")
    args = ap.parse_args(argv)
    if args.local or args.hosts is None:
        hosts, local = ["localhost"], True
    else:
        hosts = [h.strip() for h in open(args.hosts) if h.strip()]
        local = False
    with ThreadPoolExecutor(max_workers=min(len(hosts), 64)) as ex:
        while True:
            stats = dict(zip(hosts, ex.map(lambda h: query_host(h, local), hosts)))
            if not args.once:
                sys.stdout.write("\033[2J\033[H")
            print(render(stats), flush=True)
            if args.once:
                return stats
            time.sleep(args.poll_freq / 1000.0)


if __name__ == "__main__":
    main()
