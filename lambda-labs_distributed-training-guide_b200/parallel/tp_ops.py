"""Differentiable collectives of the tensor-parallel / sequence-parallel plan.

The reference expresses them as DTensor layout changes (``06-tensor-parallel/train_llm.py:84-128``):
  * ``PrepareModuleInput(Shard(1) -> Replicate())`` in front of attention / MLP and the ``lm_head`` input
    layout  -> all-gather on the sequence (backward: reduce-scatter of dX)            [N7]
  * ``RowwiseParallel(output_layouts=Shard(1))``  -> reduce-scatter on the sequence (backward: all-gather) [N8]
  * ``ColwiseParallel(output_layouts=Shard(1))`` on ``embed_tokens`` -> hidden-sharded lookup result
    exchanged into a sequence-sharded activation                                               [N9]
  * replicated (SequenceParallel) norm weights receive partial gradients that are summed over tp [N11]
Activations are flattened [T, H]; "sequence sharded" means rank r of the tp group holds the contiguous
token rows [r*T/tp, (r+1)*T/tp).

``TPGroup`` wraps the process group; the collectives go through ``torch.distributed`` (NCCL / gloo) unless
a peer-memory implementation is attached by ``parallel/peer.py``.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class TPGroup:
    def __init__(self, group, rank: int, size: int):
        self.group, self.rank, self.size = group, rank, size
        self.backend = dist.get_backend(group) if (size > 1 and dist.is_initialized()) else None

    # ---- raw collectives on [rows, cols] tensors, sharded / gathered along dim 0 ----
    def all_gather_rows(self, x: torch.Tensor) -> torch.Tensor:
        if self.size == 1:
            return x
        x = x.contiguous()
        out = torch.empty((x.shape[0] * self.size,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=self.group)
        return out

    def reduce_scatter_rows(self, x: torch.Tensor) -> torch.Tensor:
        if self.size == 1:
            return x
        x = x.contiguous()
        n = x.shape[0] // self.size
        out = torch.empty((n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        if self.backend == "nccl":
            dist.reduce_scatter_tensor(out, x, group=self.group)
        else:  # gloo: all-reduce, keep the owned rows
            tmp = x.float() if x.dtype == torch.bfloat16 else x.clone()
            dist.all_reduce(tmp, group=self.group)
            out.copy_(tmp[self.rank * n:(self.rank + 1) * n])
        return out

    def all_reduce_(self, x: torch.Tensor) -> torch.Tensor:
        if self.size == 1:
            return x
        if self.backend == "nccl" or x.dtype != torch.bfloat16:
            dist.all_reduce(x, group=self.group)
        else:
            tmp = x.float()
            dist.all_reduce(tmp, group=self.group)
            x.copy_(tmp)
        return x

    def all_to_all_rows(self, x: torch.Tensor) -> torch.Tensor:
        """Input: ``size`` equal row blocks, block j destined for rank j. Output: the blocks received."""
        if self.size == 1:
            return x
        x = x.contiguous()
        out = torch.empty_like(x)
        if self.backend == "nccl":
            dist.all_to_all_single(out, x, group=self.group)
        else:  # gloo (CPU tests): all-gather everything, keep the blocks addressed to this rank
            gathered = [torch.empty_like(x) for _ in range(self.size)]
            dist.all_gather(gathered, x, group=self.group)
            n = x.shape[0] // self.size
            for j in range(self.size):
                out[j * n:(j + 1) * n].copy_(gathered[j][self.rank * n:(self.rank + 1) * n])
        return out


class _AllGatherSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        return tp.all_gather_rows(x)

    @staticmethod
    def backward(ctx, g):
        return ctx.tp.reduce_scatter_rows(g), None


class _ReduceScatterSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        return tp.reduce_scatter_rows(x)

    @staticmethod
    def backward(ctx, g):
        return ctx.tp.all_gather_rows(g), None


class _HiddenToSeq(torch.autograd.Function):
    """[T, H/tp] (all tokens, local hidden slice) -> [T/tp, H] (local tokens, full hidden)."""

    @staticmethod
    def forward(ctx, x, tp):
        ctx.tp = tp
        T, Hl = x.shape
        fz = getattr(tp, "fused", None)
        if fz is not None and x.is_cuda:                   # strided copy-engine pulls from peer memory
            return fz.hidden_to_seq(x.contiguous())
        recv = tp.all_to_all_rows(x)                       # block j: my tokens' hidden slice j
        n = T // tp.size
        return recv.view(tp.size, n, Hl).transpose(0, 1).reshape(n, tp.size * Hl)

    @staticmethod
    def backward(ctx, g):
        tp = ctx.tp
        n, H = g.shape
        Hl = H // tp.size
        fz = getattr(tp, "fused", None)
        if fz is not None and g.is_cuda:
            return fz.seq_to_hidden(g.contiguous()), None
        send = g.view(n, tp.size, Hl).transpose(0, 1).contiguous().view(tp.size * n, Hl)
        return tp.all_to_all_rows(send), None


def all_gather_seq(x, tp: TPGroup):
    return x if tp.size == 1 else _AllGatherSeq.apply(x, tp)


def reduce_scatter_seq(x, tp: TPGroup):
    return x if tp.size == 1 else _ReduceScatterSeq.apply(x, tp)


def hidden_to_seq(x, tp: TPGroup):
    return x if tp.size == 1 else _HiddenToSeq.apply(x, tp)
