"""GEMMs with fused epilogues (``ops.fused``): gate/up + SwiGLU, down-projection dgrad + SwiGLU backward,
QKV projection + rotary embedding. Oracles: (1) the unfused kernel sequence, which the fused forwards must
reproduce bit for bit (the epilogues apply the elementwise math to the bf16-rounded projection, as a
separate kernel would), (2) fp32 PyTorch math for the values and for every gradient."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).to(torch.bfloat16)


def _rel(a, b):
    return (a.float() - b.float()).abs().max().item() / max(b.float().abs().max().item(), 1e-6)


@pytest.mark.parametrize("T,H,I", [(512, 512, 1024), (384, 256, 384), (1024, 1024, 2816), (4096, 4096, 14336)])
def test_mlp_swiglu_matches_unfused_and_fp32(T, H, I):
    from b200train import ops
    from b200train.ops import fused

    x = _rand((T, H), 1).requires_grad_()
    w_gu = _rand((2 * I, H), 2, H ** -0.5).requires_grad_()
    w_d = _rand((H, I), 3, I ** -0.5).requires_grad_()
    dy = _rand((T, H), 4)
    assert fused.ENABLED
    y = ops.mlp_swiglu(x, w_gu, w_d)
    y.backward(dy)
    got = [y.detach(), x.grad.clone(), w_gu.grad.clone(), w_d.grad.clone()]
    for t in (x, w_gu, w_d):
        t.grad = None
    # unfused kernel sequence
    y2 = ops.linear(ops.swiglu(ops.linear(x, w_gu)), w_d)
    y2.backward(dy)
    torch.cuda.synchronize()
    errs = {name: _rel(a, b) for name, a, b in zip(("y", "dx", "dw_gu", "dw_down"), got,
                                                   (y2.detach(), x.grad, w_gu.grad, w_d.grad))}
    assert torch.equal(got[0], y2.detach()), f"forward differs from the unfused kernels: {errs}"
    assert all(e < 2e-2 for e in errs.values()), errs
    if T * I > 1 << 24:
        return
    # fp32 oracle
    xf, wgf, wdf = (t.detach().float().requires_grad_() for t in (x, w_gu, w_d))
    gu = xf @ wgf.t()
    g, u = gu.chunk(2, dim=-1)
    yf = (torch.nn.functional.silu(g) * u) @ wdf.t()
    yf.backward(dy.float())
    assert _rel(got[0], yf) < 3e-2
    for name, a, b in zip(("dx", "dw_gu", "dw_down"), got[1:], (xf.grad, wgf.grad, wdf.grad)):
        assert _rel(a, b) < 3e-2, (name, _rel(a, b))


@pytest.mark.parametrize("T,H,nq,nkv,packed", [(512, 512, 4, 2, False), (384, 256, 2, 1, True), (4096, 4096, 32, 8, False)])
def test_linear_rope_matches_unfused(T, H, nq, nkv, packed):
    from b200train import ops

    D = 128
    N = (nq + 2 * nkv) * D
    x = _rand((T, H), 5).requires_grad_()
    w = _rand((N, H), 6, H ** -0.5).requires_grad_()
    dq = _rand((T, N), 7)
    seq_len = 256 if not packed else T
    cos, sin = ops.rope_tables(D, 4096, 500000.0, None, device="cuda")
    pos = None
    if packed:   # two packed documents: positions restart
        pos = torch.cat((torch.arange(100), torch.arange(T - 100))).to(device="cuda", dtype=torch.int32)
    y = ops.linear_rope(x, w, cos, sin, pos, nq + nkv, D, seq_len)
    y.backward(dq.clone())
    got = [y.detach().clone(), x.grad.clone(), w.grad.clone()]
    x.grad = w.grad = None
    y2 = ops.rope_qk(ops.linear(x, w), cos, sin, pos, nq + nkv, D, seq_len)
    y2.backward(dq.clone())
    torch.cuda.synchronize()
    assert torch.equal(got[0], y2.detach()), _rel(got[0], y2)
    assert torch.equal(got[1], x.grad) and torch.equal(got[2], w.grad)
    # fp32 oracle of the forward
    ref = ops.rope.rope_qk_ref((x.detach().float() @ w.detach().float().t()), cos, sin, pos, nq + nkv, D, seq_len)
    assert _rel(got[0], ref) < 3e-2


def test_llama_layer_with_and_without_fused_epilogues():
    """A whole model step with the fused epilogues against the same step with them switched off."""
    from b200train.models import build_model, load_config
    from b200train.ops import fused

    cfg = load_config("llama-smoke")
    ids = torch.randint(0, cfg.vocab_size, (2, 512), device="cuda")
    out = {}
    for on in (True, False):
        fused.ENABLED = on
        try:
            model = build_model(cfg, device="cuda", seed=0)
            loss = model(input_ids=ids, labels=ids).loss
            loss.backward()
            out[on] = (loss.item(), {n: p.grad.detach().float().clone() for n, p in model.named_parameters()})
        finally:
            fused.ENABLED = True
    assert abs(out[True][0] - out[False][0]) < 1e-6, (out[True][0], out[False][0])
    for n, g in out[True][1].items():
        assert _rel(g, out[False][1][n]) < 3e-2, n
