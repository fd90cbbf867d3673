#!/usr/bin/env python
"""The other BASELINE.json configurations, both arms, same timing rules as bench.py (W warm-ups, K timed
steps between barrier + synchronize, CUDA events, max over ranks):

  gpt2-ddp    GPT-2 124M, chapter 02 DDP (+ZeRO-1), seq 1024, batch 8/GPU
  llama70b-2d Llama-3.1-70B, chapter 07 2D FSDP x TP (--tp 4 on 8 GPUs), seq 4096, batch 1 per dp group
  llama405b   Llama-3.1-405B architecture truncated to --layers, chapter 05 FSDP + CPU offload + act. ckpt

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/config_bench.py \\
      --config gpt2-ddp [--impl reference]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))

CONFIGS = {
    "gpt2-ddp": dict(model="gpt2", strategy="ddp", ref_chapter="ddp", batch=8, seq=1024, kw=dict(zero1=True)),
    "llama70b-2d": dict(model="llama-3.1-70b", strategy="2d", ref_chapter="2d", batch=1, seq=4096, kw=dict(tp=4),
                        ref_args=("--tp", "4")),
    "llama405b": dict(model="llama-3.1-405b", strategy="fsdp", ref_chapter="fsdp405", batch=1, seq=4096,
                      kw=dict(cpu_offload=True, activation_checkpointing=True), ref_args=("--cpu-offload", "on")),
    "llama8b-tp": dict(model="llama-3.1-8b", strategy="tp", ref_chapter="tp", batch=16, seq=1024, kw=dict()),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200train", choices=["b200train", "reference"])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--batch", type=int, default=None)
    args = ap.parse_args()
    c = CONFIGS[args.config]
    if c["kw"].get("cpu_offload"):
        # host AdamW threads per rank (the reference uses OMP_NUM_THREADS=26 on 8 ranks per node)
        world_env = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
        os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 8) // world_env - 2))
    import torch
    import torch.distributed as dist

    from b200train.models import load_config, to_hf_dict
    from b200train.parallel import make_strategy
    from b200train.parallel.comm import init_distributed

    cfg = load_config(c["model"])
    if args.layers:
        cfg.num_hidden_layers = args.layers
    B, S = args.batch or c["batch"], c["seq"]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    tp = c["kw"].get("tp", world if c["strategy"] == "tp" else 1)
    dp = world // tp
    out = {"config": args.config, "impl": args.impl, "model": cfg.name, "layers": cfg.num_hidden_layers,
           "params": cfg.num_parameters(), "n_gpus": world, "batch_per_dp_group": B, "seq_len": S,
           "steps": args.steps, "warmup": args.warmup, "dtype": "bf16", "data": "synthetic"}
    if args.impl == "reference":
        import run_reference as rr

        extra = tuple(c.get("ref_args", ()))
        try:
            res = rr.run_reference(c["ref_chapter"], to_hf_dict(cfg), B, S, args.steps, args.warmup, extra_args=extra,
                                   dp_size=dp)
            out.update(tokens_per_s=round(res["tokens_per_s"], 1), ms_per_step=round(res["ms_per_step"], 2),
                       peak_alloc_gb=round(res["peak_alloc_gb"], 1))
        except Exception as e:
            out["unavailable"] = f"{type(e).__name__}: {str(e)[:300]}"
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps(out), flush=True)
        return
    rank, local_rank, world, device = init_distributed("cuda")
    kw = dict(c["kw"])
    if c["strategy"] == "tp":
        kw["tp"] = world
    eng = make_strategy(c["strategy"], cfg, device, lr=3e-5, seed=0, **kw)
    dp_rank = eng.data_parallel_coords()[0] if hasattr(eng, "data_parallel_coords") else rank

    def batch(i):
        g = torch.Generator().manual_seed(99 + i * dp + dp_rank)
        ids = torch.randint(0, cfg.vocab_size, (B, S), generator=g)
        return {"input_ids": ids.pin_memory(), "labels": ids.clone().pin_memory()}

    sys.path.insert(0, ROOT)
    from bench import ClockSampler, PEAK_BF16_TFLOPS, _physical_gpu_index   # the headline bench's sampler / peaks

    batches = [batch(i) for i in range(args.warmup + args.steps)]
    for i in range(args.warmup):
        loss = eng.train_step(batches[i])
    float(loss)
    sampler = ClockSampler(_physical_gpu_index(local_rank))
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    # the timed region is end to end through the public API: every step copies its pinned host batch to the
    # device and reads the loss back (4 bytes)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        loss = eng.train_step(batches[args.warmup + i])
        lv = loss.item()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    sampler.stop()
    sampler.join(timeout=2)
    t = torch.tensor([e0.elapsed_time(e1), wall_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t[0].item() / args.steps
    tokens = dp * B * S
    h2d = sum(x.numel() * x.element_size() for x in batches[0].values())
    out.update(tokens_per_s=round(tokens / ms * 1e3, 1), ms_per_step=round(ms, 2), final_loss=round(lv, 4),
               mfu_vs_measured_burst=round(tokens / ms * 1e3 * cfg.flops_per_token(S) / (world * PEAK_BF16_TFLOPS * 1e12), 4),
               e2e={"value": round(tokens * args.steps / max(t[0].item(), t[1].item()) * 1e3, 1), "unit": "tokens/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
               clocks=sampler.summary(),
               peak_alloc_gb=round(torch.cuda.max_memory_allocated() / 1e9, 1))
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
