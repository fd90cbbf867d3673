"""Attention backward, head_dim 128: whole work items against the heaviest items shared by a CTA pair
(``B200TRAIN_ATTN_BWD_SPLIT`` / ``_SPLIT_KB``; read once per process, so every arm is a subprocess). Prints one
JSON record per (shape, arm): median milliseconds of ``--iters`` runs timed with CUDA events, each preceded by an L2 flush,
and the largest difference of dq / dk / dv between the arms.

    python bench/attn_bwd_ab.py > profiles/attn_bwd_split_ab.jsonl
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, ROOT)

SHAPES = [("llama8b_s4096", 1, 4096, 32, 8), ("llama8b_s2048", 1, 2048, 32, 8), ("llama70b_s4096", 1, 4096, 64, 8),
          ("llama8b_b2_s4096", 2, 4096, 32, 8), ("llama8b_tp8_b16_s1024", 16, 1024, 4, 1),
          ("llama70b_tp4_s4096", 1, 4096, 16, 2)]


def arm(iters, out_dir):
    import torch

    from b200train.ops.attention import attn_bwd_raw, attn_fwd_raw

    D = 128
    flush = torch.empty(256 * 2 ** 20, dtype=torch.uint8, device="cuda")
    recs = []
    for name, B, S, nq, nkv in SHAPES:
        T = B * S
        g = torch.Generator(device="cuda").manual_seed(1)
        qkv = (torch.randn(T, (nq + 2 * nkv) * D, device="cuda", generator=g)).to(torch.bfloat16)
        cu = torch.arange(0, B + 1, device="cuda", dtype=torch.int32) * S
        q, k, v = qkv[:, :nq * D], qkv[:, nq * D:(nq + nkv) * D], qkv[:, (nq + nkv) * D:]
        o, lse = attn_fwd_raw(q, k, v, cu, S, nq, nkv, D, D ** -0.5)
        do = (torch.randn(T, nq * D, device="cuda", generator=g)).to(torch.bfloat16)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :nq * D], dqkv[:, nq * D:(nq + nkv) * D], dqkv[:, (nq + nkv) * D:]
        times = []
        for it in range(iters + 3):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            attn_bwd_raw(do, q, k, v, o, lse, dq, dk, dv, cu, S, nq, nkv, D, D ** -0.5)
            e1.record()
            torch.cuda.synchronize()
            if it >= 3:
                times.append(e0.elapsed_time(e1))
        times.sort()
        ms = times[len(times) // 2]
        flops = 2.5 * 4.0 * B * S * S * nq * D / 2
        torch.save(dqkv.cpu(), os.path.join(out_dir, f"{name}.pt"))
        recs.append({"shape": name, "arm": os.environ.get("B2_AB_ARM", "auto"), "ms": round(ms, 4),
                     "tflops": round(flops / ms / 1e9, 1)})
    print(json.dumps(recs))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=15)
    ap.add_argument("--arm", default=None)
    ap.add_argument("--kb", type=int, nargs="*", default=[], help="extra arms: share exactly the first n kv blocks")
    a = ap.parse_args()
    if a.arm is not None:
        return arm(a.iters, a.arm)
    import tempfile

    import torch

    outs = {}
    arms = [("whole", {"B200TRAIN_ATTN_BWD_SPLIT": "0"}), ("auto", {})]
    arms += [(f"kb{n}", {"B200TRAIN_ATTN_BWD_SPLIT_KB": str(n)}) for n in a.kb]
    for name_arm, extra in arms * 2:   # two interleaved rounds: clock drift hits every arm
        d = tempfile.mkdtemp()
        env = dict(os.environ, B2_AB_ARM=name_arm, **extra)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--arm", d, "--iters", str(a.iters)], env=env,
                           capture_output=True, text=True)
        if r.returncode:
            print(json.dumps({"arm": name_arm, "error": r.stderr[-2000:]}))
            continue
        for rec in json.loads(r.stdout.strip().splitlines()[-1]):
            print(json.dumps(rec), flush=True)
        outs[name_arm] = d
    if "whole" in outs:
        for name_arm, d in outs.items():
            if name_arm == "whole":
                continue
            for name, *_ in SHAPES:
                x, y = torch.load(os.path.join(outs["whole"], f"{name}.pt")), torch.load(os.path.join(d, f"{name}.pt"))
                print(json.dumps({"shape": name, "arm": name_arm,
                                  "max_abs_diff_vs_whole": (x.float() - y.float()).abs().max().item(),
                                  "max_abs": x.float().abs().max().item()}), flush=True)


if __name__ == "__main__":
    main()
