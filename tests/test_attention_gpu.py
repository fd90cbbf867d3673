"""tcgen05 flash attention vs an fp32 PyTorch softmax(QK^T)V reference (causal, GQA, varlen)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _qkv(T, nq, nkv, D, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(T, (nq + 2 * nkv) * D, device="cuda", generator=g)).to(torch.bfloat16)


CASES = [
    # (doc lengths, nq, nkv, D)
    ([128], 2, 2, 128),
    ([256], 4, 2, 128),
    ([1024, 1024], 8, 2, 128),
    ([300, 77, 515], 4, 1, 128),
    ([512], 4, 4, 64),
    ([130, 1000], 6, 2, 64),
    # the benchmarked shapes: Llama-3.1-8B heads (32 query : 8 key/value, head_dim 128) at 4096 and 8192 tokens
    # (32 / 64 key blocks per query block: the long-row and tail paths), GPT-2 heads at 2048
    ([4096], 32, 8, 128),
    ([8192], 32, 8, 128),
    ([2048], 12, 12, 64),
    ([3000, 1096, 4096], 8, 2, 128),
]


@pytest.mark.parametrize("lens,nq,nkv,D", CASES)
def test_attention_forward(lens, nq, nkv, D):
    from b200train.ops.attention import attention, attention_ref

    T = sum(lens)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    qkv = _qkv(T, nq, nkv, D, 1)
    with torch.no_grad():
        out = attention(qkv, cu, max(lens), nq, nkv, D)
    ref = attention_ref(qkv.float(), cu, max(lens), nq, nkv, D)
    torch.cuda.synchronize()
    err = (out.float() - ref.float()).abs().max().item()
    assert err < 3e-2, f"max abs err {err}"


@pytest.mark.parametrize("lens,nq,nkv,D", CASES)
def test_attention_backward(lens, nq, nkv, D):
    from b200train.ops.attention import attention, attention_ref

    T = sum(lens)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    qkv = _qkv(T, nq, nkv, D, 2).requires_grad_()
    g = torch.Generator(device="cuda").manual_seed(3)
    do = torch.randn(T, nq * D, device="cuda", generator=g).to(torch.bfloat16)
    out = attention(qkv, cu, max(lens), nq, nkv, D)
    out.backward(do)
    qr = qkv.detach().float().requires_grad_()
    ref = attention_ref(qr, cu, max(lens), nq, nkv, D)
    ref.backward(do.float())
    torch.cuda.synchronize()
    err = (qkv.grad.float() - qr.grad).abs().max().item()
    scale = qr.grad.abs().max().item()
    assert err < 3e-2 * max(scale, 1.0), f"max abs err {err} (scale {scale})"
