"""Drive an UNMODIFIED reference chapter script (copied to ``baseline/_ref/reference_src``) on synthetic
data and time it the same way ``bench.py`` times b200train.

Nothing in the reference source is edited. The harness only supplies what the offline box lacks and
observes the loop from outside:
  * ``--model-name`` is a local directory holding a ``config.json`` (no hub access),
  * the module-level ``_load_and_preprocess_data`` is replaced by a synthetic fixed-length token dataset
    (columns ``input_ids``, ``attention_mask``, ``labels``), because tokenizer + dataset downloads fail,
  * ``WANDB_MODE=disabled``,
  * the module-level ``LocalTimer`` is replaced by a subclass-compatible class that, in addition to the
    reference's own host-clock measurement, records CUDA events at the start of ``data`` and the end of
    ``update`` so the step is device-timed.
The model, FSDP wrapping, optimizer, loop and collectives are the reference's own code path.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import tempfile
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_SRC = HERE / "_ref" / "reference_src"

CHAPTERS = {
    "single": "01-single-gpu/train_llm.py",
    "ddp": "02-distributed-data-parallel/train_llm.py",
    "fsdp": "04-fully-sharded-data-parallel/train_llm.py",
    "fsdp405": "05-training-llama-405b/train_llm.py",
    "tp": "06-tensor-parallel/train_llm.py",
    "2d": "07-2d-parallel/train_llm.py",
}


def reference_available():
    if not REF_SRC.exists():
        return False, f"{REF_SRC} missing (run baseline/install_reference.sh)"
    return True, ""


def _load_module(path: Path, name: str):
    spec = importlib.util.spec_from_file_location(name, str(path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def run_reference(chapter: str, hf_config: dict, batch_size: int, seq_len: int, steps: int, warmup: int,
                  extra_args=(), dp_size=None, init_weights_dir=None, capture_losses=False) -> dict:
    """Run ``warmup + steps`` optimizer steps of the reference chapter; returns timing of the last
    ``steps`` (CUDA events, max over ranks) plus the end-to-end host-clock figure.

    Parity mode (``tests/test_parity_reference_gpu.py``): ``init_weights_dir`` is a directory of HF-named
    safetensors that replaces the random initialisation right after the script's own ``from_config`` call,
    ``capture_losses`` logs every step (``--log-freq 1``) and returns the per-step losses the script reports,
    with the loader's shuffling switched off so both sides see the same batches in the same order."""
    import torch
    import torch.distributed as dist

    os.environ.setdefault("WANDB_MODE", "disabled")
    os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29517")
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])

    mod = _load_module(REF_SRC / CHAPTERS[chapter], f"ref_{chapter}")

    # shared scratch dir: rank 0 creates it, everyone agrees on the path through an env-derived name
    tag = os.environ.get("MASTER_PORT", "0")
    work = Path(tempfile.gettempdir()) / f"b200train_ref_{tag}"
    model_dir = work / "model"
    if rank == 0:
        model_dir.mkdir(parents=True, exist_ok=True)
        (model_dir / "config.json").write_text(json.dumps(hf_config))
        for f in ("state.json",):
            p = work / "out" / "bench" / f
            if p.exists():
                p.unlink()

    if chapter == "fsdp405":
        # chapter 05 loads the model on rank 0 with from_pretrained(--model-name) and reads HF_HOME: give it a
        # local checkpoint directory -- random-init weights of the named architecture written once as safetensors
        # (there is no network for the real checkpoint)
        os.environ.setdefault("HF_HOME", str(work / "hf_home"))
        marker = model_dir / "weights.done"
        if rank == 0 and not marker.exists():
            from transformers import AutoConfig, AutoModelForCausalLM

            nthreads = torch.get_num_threads()
            torch.set_num_threads(max(nthreads, min(32, os.cpu_count() or 1)))
            cfg_obj = AutoConfig.from_pretrained(str(model_dir))
            m = AutoModelForCausalLM.from_config(cfg_obj, torch_dtype=torch.bfloat16)
            m.save_pretrained(str(model_dir), safe_serialization=True)
            del m
            torch.set_num_threads(nthreads)
            marker.write_text("ok")
        while not marker.exists():
            time.sleep(1.0)

    vocab = hf_config["vocab_size"]
    dp = dp_size or world            # chapters 06/07 sample over the dp coordinate only
    n_items = (warmup + steps) * batch_size * dp

    class SyntheticTokens(torch.utils.data.Dataset):
        def __len__(self):
            return n_items

        def __getitem__(self, i):
            g = torch.Generator().manual_seed(1234 + i)
            ids = torch.randint(0, vocab, (seq_len,), generator=g)
            return {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": ids.clone()}

    mod._load_and_preprocess_data = lambda args, config: SyntheticTokens()

    records = {"starts": [], "ends": [], "host_start": [], "host_end": []}
    order = []

    class EventTimer(mod.LocalTimer):
        def __init__(self, device):
            super().__init__(device)
            self._idx = len(order)
            order.append(self)
            self._dev = device

        def __enter__(self):
            if self._idx % 4 == 0 and self._dev.type == "cuda":
                ev = torch.cuda.Event(enable_timing=True)
                records["host_start"].append(time.perf_counter())
                super().__enter__()
                ev.record()
                records["starts"].append(ev)
                return self
            return super().__enter__()

        def __exit__(self, type, value, traceback):
            if self._idx % 4 == 3 and self._dev.type == "cuda" and traceback is None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                records["ends"].append(ev)
                out = super().__exit__(type, value, traceback)
                records["host_end"].append(time.perf_counter())
                return out
            return super().__exit__(type, value, traceback)

    mod.LocalTimer = EventTimer

    losses = []
    if init_weights_dir is not None:
        from safetensors.torch import load_file

        orig_from_config = mod.AutoModelForCausalLM.from_config

        def from_config_with_weights(*a, **kw):
            model = orig_from_config(*a, **kw)
            sd = {}
            for f in sorted(Path(init_weights_dir).glob("*.safetensors")):
                sd.update(load_file(str(f)))
            missing, unexpected = model.load_state_dict(sd, strict=False)
            missing = [k for k in missing if "rotary" not in k and "inv_freq" not in k]
            assert not missing and not unexpected, (missing[:4], unexpected[:4])
            return model

        class _Auto:
            from_config = staticmethod(from_config_with_weights)
            from_pretrained = staticmethod(mod.AutoModelForCausalLM.from_pretrained)

        mod.AutoModelForCausalLM = _Auto
    if capture_losses:
        orig_loader = mod.DataLoader

        def loader_in_order(*a, **kw):
            kw["shuffle"] = False
            return orig_loader(*a, **kw)

        mod.DataLoader = loader_in_order
        orig_info = mod.LOGGER.info

        def info(msg, *a, **kw):
            if isinstance(msg, dict) and "running_loss" in msg:
                losses.append(float(msg["running_loss"]))
            return orig_info(msg, *a, **kw)

        mod.LOGGER.info = info

    argv = ["train_llm.py", "-e", "bench", "-d", "synthetic", "-m", str(model_dir), "--save-dir",
            str(work / "out"), "-b", str(batch_size), "-s", str(seq_len), "--num-epochs", "1",
            "--log-freq", "1" if capture_losses else str(10 ** 9), "--ckpt-freq", str(10 ** 9), *extra_args]
    old_argv = sys.argv
    sys.argv = argv
    try:
        inner = getattr(mod.main, "__wrapped__", mod.main)  # @record wrapper -> plain function
        t0 = time.perf_counter()
        inner()
        t_end = time.perf_counter()
    finally:
        sys.argv = old_argv
    torch.cuda.synchronize()
    n = min(len(records["starts"]), len(records["ends"]))
    assert n >= warmup + steps, f"reference ran {n} steps, expected {warmup + steps}"
    dev_ms = [records["starts"][i].elapsed_time(records["ends"][i]) for i in range(warmup, warmup + steps)]
    total_dev_ms = sum(dev_ms)
    e2e_s = t_end - records["host_start"][warmup]
    t = torch.tensor([total_dev_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if dist.is_initialized() and world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_dev_ms, e2e_ms = t.tolist()
    tokens = dp * batch_size * seq_len * steps
    return {
        "ms_per_step": total_dev_ms / steps,
        "tokens_per_s": tokens / (total_dev_ms / 1e3),
        "e2e_tokens_per_s": tokens / (e2e_ms / 1e3),
        "world": world,
        "setup_s": records["host_start"][0] - t0,
        "peak_alloc_gb": torch.cuda.max_memory_allocated() / 1e9,
        "losses": losses,
    }
