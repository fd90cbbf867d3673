"""Dropout (GPT-2's three sites) with Philox masks regenerated in backward.

Elementwise kernel: keep rate, 1/(1-p) scaling, determinism under a fixed (seed, offset), backward uses the
forward's mask. Attention-probability dropout inside the attention kernels: the mask the kernels apply is
read back with ``b2_attn_dropout_mask`` and fed to an fp32 PyTorch reference, so output and all three
gradients are compared exactly (same mask), for head_dim 64 (GPT-2) and 128; plus the keep-rate statistics.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_elementwise_dropout_statistics_and_backward():
    from b200train import ops
    from b200train.ops.dropout import _Dropout

    x = torch.randn(1 << 20, device="cuda").to(torch.bfloat16).requires_grad_()
    p = 0.1
    y = _Dropout.apply(x, p, 1234, 99)
    y2 = _Dropout.apply(x, p, 1234, 99)
    y3 = _Dropout.apply(x, p, 1234, 100)
    assert torch.equal(y, y2) and not torch.equal(y, y3)          # a pure function of (seed, offset)
    keep = y != 0
    rate = keep.float().mean().item()
    assert abs(rate - (1 - p)) < 3e-3, rate                      # ~1e6 samples: sigma = 3e-4
    ref = (x.detach().float() / (1 - p)).to(torch.bfloat16)
    assert torch.equal(y[keep], ref[keep])                        # kept values scaled by 1 / (1 - p)
    g = torch.randn_like(y)
    y.backward(g)
    gref = torch.where(keep, (g.float() / (1 - p)).to(torch.bfloat16), torch.zeros_like(g))
    assert torch.equal(x.grad, gref)                              # backward regenerates the same mask
    torch.manual_seed(5)
    a = ops.dropout(x.detach(), p)
    torch.manual_seed(5)
    b = ops.dropout(x.detach(), p)
    assert torch.equal(a, b)                                      # torch.manual_seed makes runs repeatable
    assert ops.dropout(x, p, training=False) is x and ops.dropout(x, 0.0) is x


def test_host_and_device_dropout_drop_the_same_elements():
    """``csrc/cpu_dropout.cpp`` mirrors the device kernel's Philox layout: the same (seed, offset) gives the same
    output bit for bit on a CPU tensor and on a CUDA tensor (also across the AVX-512 / scalar-tail boundary)."""
    from b200train.ops.dropout import dropout

    for n in (8 * 1000, 128 * 77 + 8 * 5):
        x = torch.randn(n).to(torch.bfloat16)
        for p, philox in ((0.1, (1234, 99)), (0.5, (2 ** 62 + 12345, 2 ** 40 + 7))):
            host = dropout(x, p, philox=philox)
            dev = dropout(x.cuda(), p, philox=philox)
            assert torch.equal(host, dev.cpu()), (n, p)


def _keep_mask(nq, S, p, seed, offset):
    from b200train import _C
    from b200train.ops.dispatch import stream

    keep = torch.empty((nq, S, S), dtype=torch.uint8, device="cuda")
    _C.call("b2_attn_dropout_mask", keep.data_ptr(), nq, S, 0, p, seed, offset, stream(keep))
    return keep.bool()


@pytest.mark.parametrize("S,nq,nkv,D", [(256, 2, 2, 64), (1024, 12, 12, 64), (384, 4, 2, 128), (200, 2, 1, 64)])
def test_attention_dropout_matches_reference_with_the_same_mask(S, nq, nkv, D):
    from b200train.ops.attention import attention

    p, seed, offset = 0.1, 777, 31337
    g = torch.Generator(device="cuda").manual_seed(S + D)
    qkv = torch.randn(S, (nq + 2 * nkv) * D, device="cuda", generator=g).to(torch.bfloat16).requires_grad_()
    do = torch.randn(S, nq * D, device="cuda", generator=g).to(torch.bfloat16)
    cu = torch.tensor([0, S], dtype=torch.int32, device="cuda")
    out = attention(qkv, cu, S, nq, nkv, D, dropout_p=p, philox=(seed, offset))
    out2 = attention(qkv, cu, S, nq, nkv, D, dropout_p=p, philox=(seed, offset))
    assert torch.equal(out, out2)
    out.backward(do)
    keep = _keep_mask(nq, S, p, seed, offset)                       # [nq, S, S]
    causal = torch.ones(S, S, dtype=torch.bool, device="cuda").tril()
    rate = keep[:, causal].float().mean().item()
    p_eff = round(p * 256) / 256
    assert abs(rate - (1 - p_eff)) < 6e-3, rate
    # fp32 reference with that mask
    x = qkv.detach().float().requires_grad_()
    q = x[:, : nq * D].reshape(S, nq, D).transpose(0, 1)
    k = x[:, nq * D: (nq + nkv) * D].reshape(S, nkv, D).transpose(0, 1).repeat_interleave(nq // nkv, dim=0)
    v = x[:, (nq + nkv) * D:].reshape(S, nkv, D).transpose(0, 1).repeat_interleave(nq // nkv, dim=0)
    att = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    att = att.masked_fill(~causal, float("-inf"))
    pr = torch.softmax(att, dim=-1)
    pr = torch.where(keep, pr / (1 - p_eff), torch.zeros_like(pr))
    ref = (pr @ v).transpose(0, 1).reshape(S, nq * D)
    ref.backward(do.float())
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err < 3e-2, err
    gerr = (qkv.grad.float() - x.grad).abs().max().item()
    scale = x.grad.abs().max().item()
    assert gerr < 3e-2 * max(scale, 1.0), (gerr, scale)


def test_gpt2_trains_with_dropout():
    from b200train.models import load_config
    from b200train.parallel import make_strategy

    cfg = load_config("gpt2-tiny")
    cfg.dropout = 0.1
    torch.manual_seed(0)
    eng = make_strategy("single", cfg, "cuda", lr=1e-3, seed=0)
    ids = torch.randint(0, cfg.vocab_size, (4, 256))
    batch = {"input_ids": ids, "labels": ids}
    losses = [eng.train_step(batch).item() for _ in range(10)]
    assert all(math.isfinite(x) for x in losses) and losses[-1] < losses[0] - 0.3, losses
    # evaluation mode: no dropout, deterministic
    eng.model.eval()
    with torch.no_grad():
        a = eng.model(input_ids=ids.cuda(), labels=ids.cuda()).loss.item()
        b = eng.model(input_ids=ids.cuda(), labels=ids.cuda()).loss.item()
    assert a == b
