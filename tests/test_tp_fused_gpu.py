"""Fused tensor-parallel linears on real GPUs (``b2_gemm_ag``: sequence all-gather -> GEMM, ``b2_gemm_rs``:
GEMM -> sequence reduce-scatter; ``parallel/tp_fused.py``), forward and backward, against fp32 references built
from NCCL collectives + matmul on the same operands. Reference semantics: ``ColwiseParallel`` behind a
``Shard(1) -> Replicate`` input and ``RowwiseParallel(output_layouts=Shard(1))``
(``06-tensor-parallel/train_llm.py:89-115``). Needs at least 2 GPUs; runs on all visible ones (2, 4 or 8).
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _close(got, ref, what, tol=2e-2):
    err = (got.float() - ref).abs().max().item()
    scale = max(ref.abs().max().item(), 1.0)
    return (what, err <= tol * scale, err, scale)


def _worker(rank, world, port, ret):
    import torch.distributed as dist

    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device(f"cuda:{rank}")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from b200train.parallel.tp_fused import FusedTP, _ColwiseFused, _RowwiseFused

        fz = FusedTP(None, rank, world, dev)
        g = torch.Generator(device=dev).manual_seed(100 + rank)

        def rnd(*shape, scale=0.5):
            return (torch.randn(*shape, device=dev, generator=g) * scale).to(torch.bfloat16)

        n = 512                      # rows per rank: a multiple of 256 (two 128-row tiles per CTA pair)
        T = n * world
        checks = []
        for rep in range(2):         # twice: the double-buffered symmetric buffers and the epochs wrap around
            # ---- column parallel: y[T, N] = AG(x_local)[T, K] @ w[N, K]^T ----
            K, N = 512, 768
            x_local, w = rnd(n, K).requires_grad_(), rnd(N, K, scale=0.05).requires_grad_()
            y = _ColwiseFused.apply(x_local, w, fz)
            full = torch.empty(T, K, dtype=torch.bfloat16, device=dev)
            dist.all_gather_into_tensor(full, x_local.detach())
            checks.append(_close(y, full.float() @ w.detach().float().t(), f"colwise fwd {rep}"))
            dy = rnd(T, N)
            y.backward(dy)
            dx_full = dy.float() @ w.detach().float()            # this rank's partial sum of dX over its columns
            dist.all_reduce(dx_full)
            checks.append(_close(x_local.grad, dx_full[rank * n:(rank + 1) * n], f"colwise dx {rep}"))
            checks.append(_close(w.grad, dy.float().t() @ full.float(), f"colwise dw {rep}", tol=3e-2))
            # ---- row parallel: y_local[n, N] = RS(x[T, Kl] @ w[N, Kl]^T) ----
            Kl, N = 512, 1024
            x, w = rnd(T, Kl).requires_grad_(), rnd(N, Kl, scale=0.05).requires_grad_()
            y_local = _RowwiseFused.apply(x, w, fz)
            y_full = x.detach().float() @ w.detach().float().t()
            dist.all_reduce(y_full)
            checks.append(_close(y_local, y_full[rank * n:(rank + 1) * n], f"rowwise fwd {rep}"))
            dy_local = rnd(n, N)
            y_local.backward(dy_local)
            dy_full = torch.empty(T, N, dtype=torch.bfloat16, device=dev)
            dist.all_gather_into_tensor(dy_full, dy_local)
            checks.append(_close(x.grad, dy_full.float() @ w.detach().float(), f"rowwise dx {rep}"))
            checks.append(_close(w.grad, dy_full.float().t() @ x.detach().float(), f"rowwise dw {rep}", tol=3e-2))
        torch.cuda.synchronize()
        dist.barrier()
        ret[rank] = checks
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_fused_tp_linears_match_collective_plus_matmul():
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    if world not in (2, 4, 8):
        world = 2 if world < 4 else 4
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    for r in range(world):
        bad = [c for c in ret[r] if not c[1]]
        assert not bad, (r, bad)
