"""Packed-sequence collation (chapter 00 of the fork: ``00-rime/packed_dataset.py``).

A packed row holds several documents separated by EOS. The collator derives **position ids that
restart at 0 for every document** (``packed_dataset.py:35``); the model turns those into ``cu_seqlens``
for the variable-length attention kernel, which is what HF's FA2 path does from ``position_ids``
(``:46-48``). The reference also builds a T x T additive block-diagonal mask and then discards it
(``:27-32,45``); it is only materialised here on request (``return_mask=True``) for tests.
"""
from __future__ import annotations

import torch

EOS_ID_DEFAULT = 128_262  # the fork's extended-vocabulary EOS (``packed_dataset.py:14``)


def packed_positions(x: torch.Tensor, eos_id: int = EOS_ID_DEFAULT):
    """``x``: 1-D token row. Returns ``(position_ids [T], doc_lengths [n_docs])``. The last token is
    treated as a document end, as the reference forces ``x[-1] = eos_id`` (``packed_dataset.py:16``)
    -- without mutating the caller's tensor."""
    T = x.numel()
    is_end = x == eos_id
    is_end = is_end.clone()
    is_end[-1] = True
    ends = torch.nonzero(is_end).flatten()
    lengths = torch.diff(ends, prepend=torch.tensor([-1], device=x.device))
    starts = ends - lengths + 1
    doc_start_per_token = torch.repeat_interleave(starts, lengths)
    pos = torch.arange(T, device=x.device) - doc_start_per_token
    return pos, lengths


def block_diagonal_causal_mask(lengths: torch.Tensor, dtype=torch.bfloat16):
    """Additive mask [T, T]: 0 where attention is allowed (same document, causal), -inf elsewhere."""
    T = int(lengths.sum())
    doc = torch.repeat_interleave(torch.arange(lengths.numel()), lengths)
    same = doc[:, None] == doc[None, :]
    allowed = same & torch.ones(T, T, dtype=torch.bool).tril()
    mask = torch.zeros(T, T, dtype=dtype)
    return mask.masked_fill(~allowed, float("-inf"))


def cu_seqlens_from_positions(position_ids: torch.Tensor) -> torch.Tensor:
    flat = position_ids.reshape(-1)
    starts = torch.nonzero(flat == 0).flatten().to(torch.int32)
    return torch.cat((starts, torch.tensor([flat.numel()], dtype=torch.int32, device=flat.device)))


def pad_and_collate(features, eos_id: int = EOS_ID_DEFAULT, return_mask: bool = False):
    """Batch of packed rows -> ``input_ids``, ``labels`` (= input_ids), ``position_ids``
    (``packed_dataset.py:39-49``). Rows are equal length (the dataset is pre-packed)."""
    rows = [f["input_ids"] if torch.is_tensor(f["input_ids"]) else torch.as_tensor(f["input_ids"]) for f in features]
    ids = torch.stack(rows)
    pos, masks = [], []
    for r in rows:
        p, lengths = packed_positions(r, eos_id)
        pos.append(p)
        if return_mask:
            masks.append(block_diagonal_causal_mask(lengths))
    out = {"input_ids": ids, "labels": ids.clone(), "position_ids": torch.stack(pos)}
    if return_mask:
        out["attention_mask"] = torch.stack(masks)[:, None]
    return out
