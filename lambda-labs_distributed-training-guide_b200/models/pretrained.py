"""Loading pretrained Hugging Face checkpoints from a local directory into any strategy.

Reference behaviour: chapter 05 loads the full model on rank 0's CPU with ``from_pretrained`` and lets FSDP
broadcast it unit by unit (``05-training-llama-405b/train_llm.py:81-96,108-117``, 764 GB of host RAM on
one rank); chapter 00 loads Llama-3.2-3B and grows the vocabulary by 28,683 tokens with
``resize_token_embeddings`` (``00-rime/train_llm_01-single-gpu.py:49-66``).

Here every rank reads exactly the tensors of the unit it is filling, lazily, from the ``*.safetensors``
shards (``model.safetensors.index.json`` maps names to files), converts them to the strategy's sharding
through ``Strategy.unit_store_named`` and drops them, so peak host memory is one unit, not one model.
If the model's vocabulary is larger than the checkpoint's (added tokens), the new embedding / head rows
are set to the mean of the existing rows, which is what HF's ``resize_token_embeddings`` does by default.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Dict

import torch


class SafetensorsDir:
    """Name -> tensor access over a directory of safetensors shards (or a single ``model.safetensors``)."""

    def __init__(self, path):
        self.path = Path(path)
        index = self.path / "model.safetensors.index.json"
        if index.exists():
            self.weight_map: Dict[str, str] = json.loads(index.read_text())["weight_map"]
        else:
            files = sorted(self.path.glob("*.safetensors"))
            if not files:
                raise FileNotFoundError(f"no *.safetensors under {self.path}")
            from safetensors import safe_open

            self.weight_map = {}
            for f in files:
                with safe_open(str(f), framework="pt") as fh:
                    for k in fh.keys():
                        self.weight_map[k] = f.name
        self._open = {}

    def __contains__(self, name):
        return name in self.weight_map

    def get(self, name: str) -> torch.Tensor:
        from safetensors import safe_open

        fname = self.weight_map[name]
        if fname not in self._open:
            self._open[fname] = safe_open(str(self.path / fname), framework="pt")
        return self._open[fname].get_tensor(name)


def _fit_vocab(name: str, t: torch.Tensor, want_rows: int) -> torch.Tensor:
    if t.dim() == 2 and t.shape[0] < want_rows and ("embed_tokens" in name or "lm_head" in name or "wte" in name):
        extra = t.float().mean(0, keepdim=True).to(t.dtype).expand(want_rows - t.shape[0], -1)
        return torch.cat((t, extra), 0)
    return t


@torch.no_grad()
def load_pretrained(strategy, path, strict: bool = True) -> int:
    """Fill ``strategy``'s parameters from the checkpoint directory ``path``. Returns the number of tensors
    loaded. Works for single / DDP / FSDP (any sharding expressed through ``unit_store_named``) and for
    tensor-parallel strategies (each rank keeps its slice along ``tp_shard_dim``)."""
    src = SafetensorsDir(path)
    tp = getattr(strategy, "tpg", None)
    loaded = 0
    for i in range(strategy.num_units()):
        names = list(strategy.unit_named_full(i, "param", names_only=True))
        views = strategy.unit_named_full(i, "param")
        tensors = {}
        for n in names:
            key = n
            if key not in src and key == "lm_head.weight" and "model.embed_tokens.weight" in src:
                key = "model.embed_tokens.weight"          # tied checkpoints store the table once
            if key not in src:
                if strict:
                    raise KeyError(f"{n} not found in {path}")
                tensors[n] = views[n]
                continue
            t = src.get(key)
            want = views[n].shape
            if tp is not None and tp.size > 1:
                d = strategy.model.tp_shard_dim(n)
                if d is not None:
                    full_rows = want[0] * tp.size if d == 0 else want[0]
                    t = _fit_vocab(n, t, full_rows)
                    k = t.shape[d] // tp.size
                    t = t.narrow(d, tp.rank * k, k)
            t = _fit_vocab(n, t, want[0]) if t.dim() == 2 else t
            if tuple(t.shape) != tuple(want):
                raise ValueError(f"{n}: checkpoint shape {tuple(t.shape)} != model shape {tuple(want)}")
            tensors[n] = t
            loaded += 1
        strategy.unit_store_named(i, "param", tensors)
    strategy.barrier()
    return loaded


def save_pretrained(strategy, path, max_shard_bytes: int = 4 << 30):
    """Write HF-named safetensors (+ index) from rank 0 -- the inverse of :func:`load_pretrained`, used by
    tests and for exporting a trained model."""
    from safetensors.torch import save_file

    path = Path(path)
    full = {}
    for i in range(strategy.num_units()):
        for n, t in strategy.unit_named_full(i, "param").items():
            full[n] = t.detach().to("cpu").clone().contiguous()
    if strategy.rank == 0:
        path.mkdir(parents=True, exist_ok=True)
        save_file(full, str(path / "model.safetensors"))
        (path / "model.safetensors.index.json").write_text(json.dumps(
            {"metadata": {}, "weight_map": {k: "model.safetensors" for k in full}}))
    strategy.barrier()
