"""Training data: synthetic fixed-length tokens (offline benchmarks) and the reference's text pipeline.

``load_and_preprocess_data`` follows the reference function of the same name
(``01-single-gpu/train_llm.py:194-247``): tokenize the ``text`` column (or the first column), concatenate
everything, cut into ``seq_length`` blocks dropping the remainder, ``labels = input_ids``; ``seq_length``
falls back to the tokenizer's ``model_max_length`` and is clamped to ``min(1024, max_position_embeddings)``
when it exceeds the model's positions (``:218-220``).

Dataset names:
  ``synthetic`` / ``synthetic:<n_samples>``  random tokens of the model's vocabulary (no files needed)
  a local ``.txt`` / ``.jsonl`` / directory    read with ``datasets`` when importable, else plain Python
  anything else                                ``datasets.load_dataset(name)`` (needs a populated HF cache)
"""
from __future__ import annotations

import json
import logging
import multiprocessing
import os
from itertools import chain
from pathlib import Path

import torch
from torch.utils.data import Dataset


class SyntheticTokens(Dataset):
    """Deterministic random token rows; sample ``i`` depends only on ``(seed, i)``."""

    def __init__(self, vocab_size: int, seq_length: int, n_samples: int = 4096, seed: int = 0, eos_id=None,
                 doc_length: int = 701):
        """``eos_id``: make the rows PACKED -- every ``doc_length``-th token is ``eos_id`` (and no other token is),
        so a row holds ``seq_length / doc_length`` documents the way the fork's pre-packed dataset does
        (``00-rime/packed_dataset.py``: EOS-separated documents in 8192-token rows)."""
        self.vocab_size, self.seq_length, self.n, self.seed = vocab_size, seq_length, n_samples, seed
        self.eos_id, self.doc_length = eos_id, doc_length

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + int(i))
        if self.eos_id is None:
            ids = torch.randint(0, self.vocab_size, (self.seq_length,), generator=g)
        else:
            ids = torch.randint(0, self.vocab_size - 1, (self.seq_length,), generator=g)
            ids += (ids >= self.eos_id).long()          # skip the EOS id: it only appears as a separator
            ids[self.doc_length - 1::self.doc_length] = self.eos_id
        return {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": ids.clone()}


LOGGER = logging.getLogger(__name__)


class TokenBlocks(Dataset):
    """Fixed-length blocks cut from one long token stream."""

    def __init__(self, tokens: torch.Tensor, seq_length: int):
        n = (tokens.numel() // seq_length) * seq_length
        self.blocks = tokens[:n].view(-1, seq_length)

    def __len__(self):
        return self.blocks.shape[0]

    def __getitem__(self, i):
        ids = self.blocks[i]
        return {"input_ids": ids, "attention_mask": torch.ones_like(ids), "labels": ids.clone()}


class ByteTokenizer:
    """Offline fallback when no HF tokenizer files are available: UTF-8 bytes + EOS."""

    model_max_length = 1024
    eos_token_id = 256
    vocab_size = 257

    def __call__(self, texts):
        if isinstance(texts, str):
            texts = [texts]
        return {"input_ids": [list(t.encode("utf-8")) + [self.eos_token_id] for t in texts]}


def resolve_seq_length(requested, tokenizer_max, max_position_embeddings):
    seq_length = requested or tokenizer_max
    if seq_length > max_position_embeddings:
        seq_length = min(1024, max_position_embeddings)
    return seq_length


# built-in test presets have no tokenizer files anywhere; they (and an explicit opt-in) use the byte tokenizer
_BYTE_TOKENIZER_MODELS = ("llama-tiny", "llama-smoke", "gpt2-tiny")


def _load_tokenizer(model_name: str):
    """HF tokenizer of ``model_name`` (hub id or local directory), failing loudly when it cannot be loaded,
    as ``AutoTokenizer.from_pretrained`` does in the reference (``01-single-gpu/train_llm.py:199``).
    The 257-symbol byte tokenizer is used only for the built-in tiny presets or with
    ``B200TRAIN_BYTE_TOKENIZER=1``: silently training a 128k-vocabulary model on bytes is never wanted."""
    if model_name in _BYTE_TOKENIZER_MODELS or os.environ.get("B200TRAIN_BYTE_TOKENIZER") == "1":
        LOGGER.warning(f"using the offline byte-level tokenizer (vocab 257) for {model_name!r}")
        return ByteTokenizer()
    from transformers import AutoTokenizer

    try:
        return AutoTokenizer.from_pretrained(model_name, local_files_only=not os.environ.get("HF_ALLOW_NETWORK"))
    except Exception as e:
        raise RuntimeError(
            f"cannot load a tokenizer for {model_name!r} ({type(e).__name__}: {e}). Pass a local directory with "
            "tokenizer files as --model-name, set HF_ALLOW_NETWORK=1 to fetch it from the hub, or set "
            "B200TRAIN_BYTE_TOKENIZER=1 to train on raw bytes on purpose.") from e


def _read_local_texts(path: Path):
    files = sorted(path.rglob("*")) if path.is_dir() else [path]
    for f in files:
        if f.suffix == ".txt":
            yield f.read_text()
        elif f.suffix in (".jsonl", ".json"):
            for line in f.read_text().splitlines():
                if line.strip():
                    row = json.loads(line)
                    yield row["text"] if "text" in row else next(iter(row.values()))


def _cache_dir() -> Path:
    root = os.environ.get("B200TRAIN_CACHE") or os.environ.get("HF_HOME") or os.path.join(os.getcwd(), "tmp")
    d = Path(root) / "b200train-tokenized"
    d.mkdir(parents=True, exist_ok=True)
    return d


def _tokenize_local(path: Path, tokenizer, model_name: str, batch: int = 512) -> torch.Tensor:
    """One long token stream from local ``.txt`` / ``.jsonl`` files: texts go to the tokenizer in batches (the fast
    tokenizers parallelise over a batch) and the stream is cached on disk keyed by the files' names, sizes and
    modification times and the tokenizer, as ``datasets.map(load_from_cache_file=True)`` does for hub datasets
    (``01-single-gpu/train_llm.py:204-247``). The callers run this under ``rank0_first``: one rank fills the cache,
    the others read it."""
    import hashlib

    files = sorted(f for f in (path.rglob("*") if path.is_dir() else [path]) if f.is_file())
    h = hashlib.sha256(f"{model_name}|{type(tokenizer).__name__}|{getattr(tokenizer, 'vocab_size', 0)}".encode())
    for f in files:
        st = f.stat()
        h.update(f"{f.resolve()}|{st.st_size}|{st.st_mtime_ns}".encode())
    cache = _cache_dir() / f"{h.hexdigest()[:32]}.pt"
    if cache.exists():
        try:
            return torch.load(cache, weights_only=True)
        except Exception:   # a torn or foreign file: rebuild it
            pass
    chunks, pending = [], []

    def flush():
        if pending:
            for row in tokenizer(pending)["input_ids"]:
                chunks.append(torch.as_tensor(row, dtype=torch.long))
            pending.clear()

    for text in _read_local_texts(path):
        pending.append(text)
        if len(pending) >= batch:
            flush()
    flush()
    ids = torch.cat(chunks) if chunks else torch.zeros(0, dtype=torch.long)
    tmp = cache.with_name(cache.name + f".tmp{os.getpid()}")
    torch.save(ids, tmp)
    os.replace(tmp, cache)
    return ids


def group_texts(examples: dict, seq_length: int) -> dict:
    """Concatenate then chunk (the reference's inner ``group_texts``, ``01-.../train_llm.py:223-236``)."""
    concatenated = {k: list(chain(*examples[k])) for k in examples.keys()}
    total_length = len(concatenated[list(examples.keys())[0]])
    if total_length > seq_length:
        total_length = (total_length // seq_length) * seq_length
    result = {k: [t[i:i + seq_length] for i in range(0, total_length, seq_length)] for k, t in concatenated.items()}
    result["labels"] = result["input_ids"].copy()
    return result


def load_and_preprocess_data(dataset_name: str, model_name: str, config, seq_length=None, seed: int = 0,
                             packed_eos_id=None):
    """Returns a map-style dataset of ``{"input_ids", "attention_mask", "labels"}`` rows. ``packed_eos_id``: the
    synthetic rows are packed documents separated by that token (chapter 00 / ``--packed``)."""
    if dataset_name.startswith("synthetic"):
        n = int(dataset_name.split(":")[1]) if ":" in dataset_name else 4096
        sl = resolve_seq_length(seq_length, 1024, config.max_position_embeddings)
        return SyntheticTokens(config.vocab_size, sl, n, seed, eos_id=packed_eos_id)
    tokenizer = _load_tokenizer(model_name)
    sl = resolve_seq_length(seq_length, min(getattr(tokenizer, "model_max_length", 1024), 1 << 20),
                            config.max_position_embeddings)
    p = Path(dataset_name)
    if p.exists():
        return TokenBlocks(_tokenize_local(p, tokenizer, model_name), sl)
    import datasets  # hub / cache datasets, exactly the reference flow

    data = datasets.load_dataset(dataset_name, trust_remote_code=True)
    column_names = data["train"].column_names
    text_column = "text" if "text" in column_names else column_names[0]
    tokenized = data.map(lambda ex: tokenizer(ex[text_column]), batched=True, remove_columns=column_names,
                         num_proc=multiprocessing.cpu_count(), load_from_cache_file=True,
                         desc="Running tokenizer on dataset")
    lm = tokenized.map(lambda ex: group_texts(ex, sl), batched=True, num_proc=multiprocessing.cpu_count(),
                       load_from_cache_file=True, desc=f"Grouping texts in chunks of {sl}")
    return lm["train"].with_format("torch")
