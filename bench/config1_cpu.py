#!/usr/bin/env python
"""BASELINE.json configuration 1: GPT-2 124M, chapter-01 single process ON THE CPU, random init, synthetic
1024-token rows (plumbing only, no GPU). Both arms, K timed optimizer steps after W warm-ups, host clock.

  python bench/config1_cpu.py [--impl reference] [--steps 3] [--warmup 1] [--threads 16]

Reference arm: ``01-single-gpu/train_llm.py`` hard-codes ``torch.device("cuda")`` (line 40) and reads
``torch.cuda.memory_stats`` in its log record, so it cannot run on a CPU-only machine as it stands. The
harness leaves the source untouched and substitutes, in the loaded module only, a ``torch`` proxy whose
``device("cuda")`` answers ``cpu`` and a ``get_mem_stats`` that returns zeros; model, optimizer, loop and
timers are the script's own. fp32 on both sides (bf16 matmuls on CPU are an order of magnitude slower)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))


def run_ours(args):
    import torch

    from b200train.models import load_config
    from b200train.parallel import make_strategy

    torch.set_num_threads(args.threads)
    cfg = load_config("gpt2")
    eng = make_strategy("single", cfg, "cpu", lr=3e-5, seed=0, dtype=torch.bfloat16 if args.dtype == "bf16" else torch.float32)
    g = torch.Generator().manual_seed(0)
    batches = [torch.randint(0, cfg.vocab_size, (1, 1024), generator=g) for _ in range(args.warmup + args.steps)]
    for i in range(args.warmup):
        eng.train_step({"input_ids": batches[i], "labels": batches[i]})
    t0 = time.perf_counter()
    for i in range(args.steps):
        loss = eng.train_step({"input_ids": batches[args.warmup + i], "labels": batches[args.warmup + i]}).item()
    dt = time.perf_counter() - t0
    return {"tokens_per_s": round(1024 * args.steps / dt, 1), "s_per_step": round(dt / args.steps, 3),
            "final_loss": round(loss, 4)}


def run_reference(args):
    import types

    import torch

    import run_reference as rr
    from b200train.models import load_config, to_hf_dict

    torch.set_num_threads(args.threads)
    ok, why = rr.reference_available()
    if not ok:
        return {"unavailable": why}

    real_device = torch.device

    class TorchProxy(types.ModuleType):
        def __getattr__(self, k):
            return getattr(torch, k)

        @staticmethod
        def device(*a, **kw):
            if a and isinstance(a[0], str) and a[0].startswith("cuda"):
                return real_device("cpu")
            return real_device(*a, **kw)

    proxy = TorchProxy("torch")
    orig_load = rr._load_module

    def load_patched(path, name):
        mod = orig_load(path, name)
        mod.torch = proxy
        mod.get_mem_stats = lambda device: {k: 0.0 for k in ("total_gb", "curr_alloc_gb", "peak_alloc_gb",
                                                             "curr_resv_gb", "peak_resv_gb")}
        return mod

    rr._load_module = load_patched
    cfg = load_config("gpt2")
    hf = to_hf_dict(cfg)
    hf["torch_dtype"] = "float32"
    times = []
    orig_timer = None
    # the runner's CUDA-event bookkeeping is bypassed on CPU: time the whole loop from outside
    t0 = time.perf_counter()
    try:
        rr.torch = torch
        import run_reference

        # run_reference synchronises CUDA at the end; on a CPU box take the host clock around main() instead
        mod = rr._load_module(rr.REF_SRC / rr.CHAPTERS["single"], "ref_single_cpu")
        import tempfile
        from pathlib import Path

        work = Path(tempfile.mkdtemp(prefix="b200train_cfg1_"))
        (work / "model").mkdir()
        (work / "model" / "config.json").write_text(json.dumps(hf))
        n_items = args.warmup + args.steps

        class Synth(torch.utils.data.Dataset):
            def __len__(self):
                return n_items

            def __getitem__(self, i):
                g = torch.Generator().manual_seed(1234 + i)
                ids = torch.randint(0, hf["vocab_size"], (1024,), generator=g)
                return {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": ids.clone()}

        mod._load_and_preprocess_data = lambda a, c: Synth()
        stamps = []

        class T(mod.LocalTimer):
            def __enter__(self):
                stamps.append(time.perf_counter())
                return super().__enter__()

        mod.LocalTimer = T
        os.environ["WANDB_MODE"] = "disabled"
        argv = sys.argv
        sys.argv = ["train_llm.py", "-e", "cfg1", "-d", "synthetic", "-m", str(work / "model"), "--save-dir",
                    str(work / "out"), "-b", "1", "-s", "1024", "--num-epochs", "1", "--log-freq", str(10 ** 9),
                    "--ckpt-freq", str(10 ** 9)]
        try:
            getattr(mod.main, "__wrapped__", mod.main)()
        finally:
            sys.argv = argv
        t_end = time.perf_counter()
        # four timers per step: the first stamp of step k is stamps[4 * k]
        t_start = stamps[4 * args.warmup]
        dt = t_end - t_start
        return {"tokens_per_s": round(1024 * args.steps / dt, 1), "s_per_step": round(dt / args.steps, 3)}
    except Exception as e:
        return {"unavailable": f"{type(e).__name__}: {str(e)[:300]}"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200train", choices=["b200train", "reference"])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"], help="the reference script is bf16 (01-single-gpu/train_llm.py:41)")
    ap.add_argument("--threads", type=int, default=min(16, os.cpu_count() or 1))
    a = ap.parse_args()
    out = {"config": "gpt2-124m ch01 single process on CPU, seq 1024, batch 1", "dtype": a.dtype if a.impl == "b200train" else "bf16 (script default)", "impl": a.impl, "steps": a.steps,
           "warmup": a.warmup, "threads": a.threads}
    out.update(run_reference(a) if a.impl == "reference" else run_ours(a))
    print(json.dumps(out), flush=True)
