"""The shared training driver behind every chapter entry point.

One ``main()`` reproduces the loop all reference chapters copy (``01-single-gpu/train_llm.py:25-191`` and
its per-chapter deltas): parse flags -> init distributed -> build + shard the model -> data ->
optimizer / cosine schedule -> resume from ``state.json`` -> metrics sink -> epoch / step loop with the
``data / forward / backward / update`` timers -> log every ``--log-freq`` -> checkpoint every
``--ckpt-freq``. The parallel strategy is chosen with ``--strategy`` (the chapter scripts pin it).

Flags keep the reference's names and defaults. Additions are optional:
``--strategy``, ``--device``, ``--num-workers``, ``--prefetch-factor``, ``--wandb-mode``, ``--timers``,
``--activation-checkpointing``, ``--grad-accum-steps``, ``--deterministic``, ``--zero1``, ``--max-steps``,
``--sharding-strategy`` / ``--replicas`` (FSDP chapters).
"""
from __future__ import annotations

import argparse
import json
import os
from contextlib import contextmanager, nullcontext
from functools import partial
from pathlib import Path

import torch
import torch.distributed as dist

from .. import ckpt
from ..ckpt import dcp
from ..data import load_and_preprocess_data, make_dataloader, pad_and_collate, set_epoch
from ..models import load_config
from ..obs import LOGGER, MetricsSink, build_record, get_mem_stats, make_timers, setup_logging
from ..parallel import make_strategy
from ..parallel.comm import init_distributed

try:  # error capture to $TORCHELASTIC_ERROR_FILE (reference: @record, 02-.../train_llm.py:32)
    from torch.distributed.elastic.multiprocessing.errors import record
except Exception:  # pragma: no cover
    def record(fn):
        return fn


def get_parser(chapter: str = "generic") -> argparse.ArgumentParser:
    p = argparse.ArgumentParser()
    req = chapter != "00-rime"
    p.add_argument("-e", "--experiment-name", default=None, required=True)
    p.add_argument("-d", "--dataset-name", default=None if req else "synthetic", required=req)
    p.add_argument("-m", "--model-name", default=None if req else "llama-3.2-3b", required=req)
    p.add_argument("--save-dir", default="../outputs")
    p.add_argument("--seed", default=0, type=int)
    p.add_argument("--num-epochs", default=1 if chapter == "00-rime" else 100, type=int)
    if chapter != "deepspeed":
        p.add_argument("--lr", default=3e-5, type=float)
        p.add_argument("-b", "--batch-size", default=1, type=int)
    p.add_argument("--log-freq", default=50 if chapter == "00-rime" else 100, type=int)
    p.add_argument("--ckpt-freq", default=500, type=int)
    p.add_argument("-s", "--seq-length", type=int,
                   default={"06-tensor-parallel": None, "07-2d-parallel": None, "00-rime": 8192}.get(chapter, 1024))
    if chapter in ("04-fsdp", "generic"):
        p.add_argument("--numel-to-wrap", default=100_000_000, type=int,
                       help="Only applies FSDP to modules with numel > this value.")
    if chapter in ("04-fsdp", "05-405b", "generic"):
        p.add_argument("--cpu-offload", default="on" if chapter == "05-405b" else "off", choices=["on", "off"])
        # the strategies tabulated in 04-fully-sharded-data-parallel/README.md:146-151 (the reference
        # hard-codes FULL_SHARD)
        p.add_argument("--sharding-strategy", default="full_shard",
                       choices=["full_shard", "shard_grad_op", "no_shard", "hybrid_shard"])
        p.add_argument("--replicas", default=None, type=int,
                       help="hybrid_shard: number of replica groups (default: one per node)")
        p.add_argument("--sync-module-states", default="off", choices=["on", "off"],
                       help="broadcast every freshly initialised unit from rank 0 (the reference's "
                            "sync_module_states=True); off: every rank initialises from the same seed, no traffic")
    if chapter in ("07-2d-parallel", "generic"):
        p.add_argument("--tp", default=8 if chapter == "07-2d-parallel" else 1, type=int)
    if chapter == "deepspeed":
        p.add_argument("--local_rank", type=int, default=None)
        p.add_argument("--deepspeed", action="store_true")
        p.add_argument("--deepspeed_config", default=None)
    # ---- additions ----
    p.add_argument("--strategy", default=None, choices=["single", "ddp", "fsdp", "tp", "2d"])
    p.add_argument("--device", default=None, choices=["cuda", "cpu"])
    p.add_argument("--num-workers", type=int,
                   default=1 if chapter in ("05-405b", "06-tensor-parallel", "07-2d-parallel") else 0)
    p.add_argument("--prefetch-factor", default=2, type=int)
    p.add_argument("--wandb-mode", default="rank0", choices=["rank0", "local_rank0", "all", "off"])
    p.add_argument("--timers", default="local", choices=["local", "events"],
                   help="local = the reference's synchronising host timers; events = CUDA events (phases may overlap)")
    p.add_argument("--time-waiting", action="store_true", help="time the barrier skew ('waiting' phase)")
    p.add_argument("--activation-checkpointing", default=None, choices=["on", "off"])
    p.add_argument("--grad-accum-steps", default=1, type=int)
    p.add_argument("--zero1", default=None, choices=["on", "off"], help="shard optimizer state (DDP)")
    p.add_argument("--fuse-optimizer", default="auto", choices=["auto", "kernel", "stream", "off"],
                   help="FSDP: 'kernel' reduces a unit's gradient and applies AdamW to the owned shard in one kernel "
                        "inside backward (the 'update' phase then only steps the schedule); 'stream' queues a separate "
                        "AdamW launch behind the reduce-scatter; 'off' updates after backward like the reference; 'auto' picks "
                        "'kernel' up to 4 ranks and 'stream' from 8")
    p.add_argument("--deterministic", action="store_true")
    p.add_argument("--progress-bar", choices=["on", "off"],
                   default="off" if chapter in ("05-405b", "06-tensor-parallel", "07-2d-parallel") else "on",
                   help="tqdm bar over the batches of an epoch on rank 0")
    p.add_argument("--ce-chunk-rows", type=int, default=2048 if chapter == "00-rime" else 0,
                   help="LM head + cross entropy in chunks of this many token rows: the logits of one chunk at a time, "
                        "recomputed in backward (0: all rows at once, fastest; the fork's chapter 00 uses Liger's "
                        "fused linear cross entropy for the same purpose)")
    p.add_argument("--packed", action="store_true", help="packed rows: derive position ids per document")
    p.add_argument("--eos-id", type=int, default=None,
                   help="document separator of packed rows (default: 128262, the fork's extended-vocabulary EOS, "
                        "when the vocabulary has it, else the last id of the vocabulary)")
    p.add_argument("--max-steps", default=None, type=int, help="stop after this many optimizer steps")
    p.add_argument("--pretrained", default=None,
                   help="local directory with *.safetensors to start from (reference: from_pretrained, ch05 / ch00)")
    p.add_argument("--add-tokens", default=28_683 if chapter == "00-rime" else 0, type=int,
                   help="grow the vocabulary by this many tokens (resize_token_embeddings, ch00)")
    return p


@contextmanager
def rank_ordered(should_go_first: bool):
    """Run the body first on the chosen ranks, then on the others
    (``rank0_first`` / ``rank_ordered``: ``02-.../train_llm.py:279-287``, ``05-.../train_llm.py:366-373``)."""
    active = dist.is_available() and dist.is_initialized()
    if should_go_first:
        yield
    if active:
        dist.barrier()
    if not should_go_first:
        yield
    if active:
        dist.barrier()


def rank0_first():
    rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0
    return rank_ordered(rank == 0)


CHAPTER_STRATEGY = {"01-single-gpu": "single", "00-rime": "single", "02-ddp": "ddp", "04-fsdp": "fsdp",
                    "05-405b": "fsdp", "06-tensor-parallel": "tp", "07-2d-parallel": "2d", "deepspeed": "fsdp",
                    "generic": "single"}


@record
def main(chapter: str = "generic", argv=None):
    args = get_parser(chapter).parse_args(argv)
    strategy_name = args.strategy or CHAPTER_STRATEGY[chapter]
    if chapter == "deepspeed":
        # the DeepSpeed variant takes lr / micro-batch from its json (DS/train_llm.py:68-77, ds_config.json)
        ds_cfg = json.loads(Path(args.deepspeed_config).read_text()) if args.deepspeed_config else {}
        args.batch_size = ds_cfg.get("train_micro_batch_size_per_gpu", 1)
        args.lr = ds_cfg.get("optimizer", {}).get("params", {}).get("lr", 3e-5)
        zero = ds_cfg.get("zero_optimization", {})
        off = zero.get("offload_param") or {}   # a dict with a "device", or the reference's plain `false`
        args.cpu_offload = "on" if (off is True or (isinstance(off, dict) and off.get("device", "none") == "cpu")) \
            else "off"
        # ZeRO stage -> sharding: 3 = parameters + gradients + optimizer state (full shard), 2 = gradients +
        # optimizer state (parameters stay gathered), 1 = optimizer state only (DDP + ZeRO-1), 0 = plain DDP
        stage = int(zero.get("stage", 3))
        if args.strategy is None and stage < 3:
            if stage == 2:
                args.sharding_strategy = "shard_grad_op"
            else:
                strategy_name = "ddp"
                args.zero1 = "on" if stage == 1 else "off"
        if args.local_rank is not None:
            os.environ.setdefault("LOCAL_RANK", str(args.local_rank))
    rank, local_rank, world_size, device = init_distributed(args.device)
    setup_logging(rank)
    LOGGER.info(os.environ)
    LOGGER.info(args)
    LOGGER.info(f"local_rank={local_rank} rank={rank} world size={world_size}")
    if args.deterministic:
        ckpt.set_determinism(args.seed)
    torch.manual_seed(args.seed)
    if args.ce_chunk_rows:
        os.environ["B200TRAIN_CE_CHUNK_ROWS"] = str(args.ce_chunk_rows)

    # node-local vs shared HF cache decides who goes first (02-.../README.md:377-408)
    first = (local_rank == 0) if os.environ.get("B200TRAIN_NODE_LOCAL_CACHE") else (rank == 0)
    with rank_ordered(first):
        config = load_config(args.model_name)
    if args.add_tokens and (args.pretrained or chapter != "00-rime"):
        config.vocab_size += args.add_tokens
    LOGGER.info(f"{config.num_parameters()} model parameters")
    LOGGER.info(f"Before sharding: {get_mem_stats(device)}")

    kw = dict(lr=args.lr, seed=args.seed)
    if device.type == "cpu":
        kw["dtype"] = torch.bfloat16 if os.environ.get("B200TRAIN_CPU_BF16") else torch.float32
    act = args.activation_checkpointing or ("on" if chapter == "05-405b" else "off")
    sharding = getattr(args, "sharding_strategy", "full_shard") if strategy_name == "fsdp" else "full_shard"
    if sharding == "no_shard":          # replicated parameters, gradients and optimizer state = plain DDP
        strategy_name = "ddp"
        args.zero1 = "off"
    if strategy_name == "fsdp":
        kw.update(cpu_offload=getattr(args, "cpu_offload", "off") == "on", activation_checkpointing=act == "on",
                  numel_to_wrap=getattr(args, "numel_to_wrap", 0),
                  sync_module_states=getattr(args, "sync_module_states", "off") == "on")
        if device.type == "cuda" and args.grad_accum_steps <= 1:
            kw.update(fuse_optimizer=False if args.fuse_optimizer == "off" else args.fuse_optimizer)
        if sharding == "shard_grad_op":  # parameters stay gathered between forward and backward (ZeRO-2)
            kw.update(reshard_after_forward=False)
        elif sharding == "hybrid_shard":
            strategy_name = "hsdp"
            kw.update(replicas=args.replicas)
    elif strategy_name == "ddp":
        kw.update(zero1=(args.zero1 or "on") == "on")
    elif strategy_name in ("tp", "2d"):
        tp = getattr(args, "tp", None)
        if strategy_name == "tp" and not (tp and tp > 1):
            # chapter 06: TP spans a node, DP spans nodes (06-tensor-parallel/train_llm.py:41-56)
            per_node = torch.cuda.device_count() if device.type == "cuda" else world_size
            tp = min(world_size, max(per_node, 1))
        kw.update(tp=tp, activation_checkpointing=act == "on")
    engine = make_strategy(strategy_name, config, device, **kw)
    if chapter == "deepspeed" and ds_cfg.get("scheduler", {}).get("type") == "WarmupCosineLR":
        # the engine builds the schedule named in the json (DS/train_llm.py:68-73, ds_config.json:9-16)
        from ..utils.schedule import make_warmup_cosine

        engine.lr_scheduler = make_warmup_cosine(engine.optimizer, **ds_cfg["scheduler"].get("params", {}))
    if args.pretrained and not ckpt.has_checkpoint(Path(args.save_dir) / args.experiment_name):
        from ..models.pretrained import load_pretrained

        n_loaded = load_pretrained(engine, args.pretrained)
        LOGGER.info(f"loaded {n_loaded} pretrained tensors from {args.pretrained}")
    LOGGER.info(f"After sharding: {get_mem_stats(device)}")

    packed = args.packed or chapter == "00-rime"
    eos_id = None
    if packed:
        from ..data.packed import EOS_ID_DEFAULT

        eos_id = args.eos_id if args.eos_id is not None else (
            EOS_ID_DEFAULT if config.vocab_size > EOS_ID_DEFAULT else config.vocab_size - 1)
        LOGGER.info(f"packed rows: documents end at token id {eos_id}")
    with rank_ordered(first):
        train_data = load_and_preprocess_data(args.dataset_name, args.model_name, config, args.seq_length,
                                              seed=args.seed, packed_eos_id=eos_id)
    seq_length = len(train_data[0]["input_ids"])
    LOGGER.info(f"{len(train_data)} training samples")
    if hasattr(engine, "data_parallel_coords"):
        dp_rank, dp_size = engine.data_parallel_coords()
    else:
        dp_rank, dp_size = rank, world_size
    dataloader = make_dataloader(train_data, args.batch_size, dp_rank=dp_rank, dp_size=dp_size, seed=args.seed,
                                 num_workers=args.num_workers, prefetch_factor=args.prefetch_factor,
                                 collate_fn=partial(pad_and_collate, eos_id=eos_id) if packed else None,
                                 pin_memory=device.type == "cuda", distributed=strategy_name != "single")
    LOGGER.info(f"{len(dataloader)} batches per epoch")

    exp_dir = Path(args.save_dir) / args.experiment_name
    sharded_ckpt = strategy_name in ("fsdp", "hsdp", "tp", "2d") or (strategy_name == "ddp" and kw.get("zero1"))

    # ---- resume ----
    state = ckpt.new_state()
    resumed = False
    if ckpt.has_checkpoint(exp_dir):
        if sharded_ckpt:
            dcp.load_sharded(exp_dir, engine)
            engine.lr_scheduler.load_state_dict(torch.load(exp_dir / "lr_scheduler.pt", weights_only=True))
            state = ckpt.load_state(exp_dir)
        else:
            blob = ckpt.load_full(exp_dir, device)
            engine.load_model_state_dict(blob["model"])
            if "optimizer" in blob:
                engine.optimizer.load_state_dict(blob["optimizer"])
            if "lr_scheduler" in blob:
                engine.lr_scheduler.load_state_dict(blob["lr_scheduler"])
            state = blob["state"]
        if args.deterministic:
            ckpt.load_rng(exp_dir, rank)
        resumed = True
    LOGGER.info(f"Resumed={resumed} | {state}")
    engine.barrier()

    # shared vs node-local experiment dir decides who creates it (04-.../train_llm.py:158-163)
    try:
        is_mount = exp_dir.is_mount()
    except Exception:
        is_mount = False
    if (is_mount and rank == 0) or (not is_mount and local_rank == 0):
        LOGGER.info("Creating experiment root directory")
        exp_dir.mkdir(parents=True, exist_ok=True)
    engine.barrier()
    if strategy_name != "single":
        (exp_dir / f"rank-{rank}").mkdir(parents=True, exist_ok=True)

    sink = MetricsSink(exp_dir, rank, local_rank, args.experiment_name, resumed, mode=args.wandb_mode,
                       config={"args": vars(args), "training_data_size": len(train_data),
                               "num_batches": len(dataloader), "world_size": world_size})
    timers = make_timers(device, args.timers, args.time_waiting)
    tok_per_step = engine.tokens_per_step(args.batch_size, seq_length) * args.grad_accum_steps
    steps_done = 0
    stop = False

    for state["epoch"] in range(state["epoch"], args.num_epochs):
        LOGGER.info(f"Begin epoch {state['epoch']} at step {state['epoch_step']}")
        set_epoch(dataloader, state["epoch"])
        batches = iter(dataloader)
        n_batches = len(dataloader)
        micro = 0
        # progress bar on rank 0 (01-single-gpu/train_llm.py:120, 02-.../train_llm.py:138); the large-scale chapters
        # switch it off, as the reference does (05-training-llama-405b/train_llm.py:217)
        bar = None
        if rank == 0 and args.progress_bar == "on":
            try:
                import tqdm

                bar = tqdm.tqdm(total=n_batches, initial=0, desc=f"epoch {state['epoch']}")
            except ImportError:
                bar = None
        for i_step in range(n_batches):
            if bar is not None:
                bar.update(1)
            with timers["data"], torch.no_grad():
                batch = next(batches)
                batch = engine.to_device(batch)
            if i_step < state["epoch_step"]:
                continue  # resume: replay the sampler up to the saved position (01-.../train_llm.py:135-137)
            if "waiting" in timers:
                with timers["waiting"]:
                    engine.barrier()
            micro += 1
            boundary = micro % args.grad_accum_steps == 0
            sync_ctx = nullcontext() if (boundary or not hasattr(engine, "no_sync")) else engine.no_sync()
            with timers["forward"]:
                outputs = engine.forward(batch)
            with timers["backward"], sync_ctx:
                if args.grad_accum_steps == 1 or micro % args.grad_accum_steps == 1:
                    engine.zero_grad()
                loss = outputs.loss / args.grad_accum_steps if args.grad_accum_steps > 1 else outputs.loss
                engine.backward(loss)
            state["epoch_step"] += 1
            state["running_loss"] += outputs.loss.item() / args.grad_accum_steps
            if not boundary:
                continue
            with timers["update"]:
                engine.optimizer_step()
            state["global_step"] += 1
            steps_done += 1

            if state["global_step"] % args.log_freq == 0:
                info = build_record(state, engine.lr_scheduler.get_last_lr()[0], args.log_freq, n_batches, i_step,
                                    tok_per_step, timers, device)
                LOGGER.info(info)
                sink.log(info, step=state["global_step"])
                if device.type == "cuda":
                    torch.cuda.reset_peak_memory_stats(device)
                state["running_loss"] = 0
                for t in timers.values():
                    t.reset()

            if state["global_step"] % args.ckpt_freq == 0:
                save_checkpoint(engine, exp_dir, state, sharded_ckpt, rank, args.deterministic)

            if args.max_steps is not None and steps_done >= args.max_steps:
                stop = True
                break
        if bar is not None:
            bar.close()
        if stop:
            break
        state["epoch_step"] = 0
    sink.close()
    return state


def save_checkpoint(engine, exp_dir, state, sharded: bool, rank: int, with_rng: bool = False):
    if sharded:
        dcp.save_sharded(exp_dir, engine, state=state, lr_scheduler=engine.lr_scheduler)
    else:
        if rank == 0:
            ckpt.save_full(exp_dir, engine.model_state_dict(), engine.optimizer.state_dict(), engine.lr_scheduler,
                           state)
        engine.barrier()
    if with_rng:
        ckpt.save_rng(exp_dir, rank)


if __name__ == "__main__":
    main()
