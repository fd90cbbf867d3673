"""tcgen05 GEMM vs an fp32 PyTorch reference of the same product (all operand layouts, both CTA-group
modes, both tile widths, ragged shapes, epilogues)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

CONFIGS = [1 | (256 << 8), 1 | (128 << 8), 2 | (256 << 8), 2 | (128 << 8)]
SHAPES = [(128, 256, 64), (256, 512, 256), (384, 640, 320), (1000, 776, 520), (2048, 4096, 1024)]


def _mk(rows, cols, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(rows, cols, device="cuda", generator=g) * 0.5).to(torch.bfloat16)


def _check(out, ref, K):
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-2 * max(scale, 1.0), f"max abs err {err} (ref scale {scale})"


@pytest.mark.parametrize("config", CONFIGS)
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("layout", ["kk", "kmn", "mnmn", "mnk"])
def test_gemm_layouts(config, shape, layout):
    from b200train.ops import gemm

    M, N, K = shape
    a_mn = layout in ("mnmn", "mnk")
    b_mn = layout in ("kmn", "mnmn")
    if a_mn and M % 64:
        M = (M + 63) // 64 * 64
    if b_mn and N % 64:
        N = (N + 63) // 64 * 64
    A = _mk(M, K, 1)
    B = _mk(N, K, 2)
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    out = gemm(a, b, a_mn=a_mn, b_mn=b_mn, config=config)
    torch.cuda.synchronize()
    _check(out, A.float() @ B.float().t(), K)


@pytest.mark.parametrize("config", [1 | (256 << 8), 2 | (256 << 8)])
def test_gemm_epilogues(config):
    from b200train.ops import gemm
    from b200train.ops.activation import gelu_new_ref

    M, N, K = 512, 768, 256
    A, B = _mk(M, K, 3), _mk(N, K, 4)
    bias = _mk(1, N, 5)[0]
    ref = A.float() @ B.float().t()
    _check(gemm(A, B, epilogue=1, bias=bias, config=config), ref + bias.float(), K)
    _check(gemm(A, B, epilogue=2, bias=bias, config=config), gelu_new_ref(ref + bias.float()), K)
    C = _mk(M, N, 6)
    C0 = C.float().clone()
    gemm(A, B, out=C, epilogue=3, config=config)
    _check(C, ref + C0, K)


def test_gemm_strided_views():
    """Operands and outputs that are column slices of wider buffers (fused QKV / padded logits)."""
    from b200train.ops import gemm

    M, N, K = 256, 320, 192
    Abig, Bbig = _mk(M, K + 64, 7), _mk(N, K + 128, 8)
    A, B = Abig[:, 64:], Bbig[:, :K]
    Cbig = torch.zeros(M, N + 64, device="cuda", dtype=torch.bfloat16)
    gemm(A, B, out=Cbig[:, :N])
    _check(Cbig[:, :N], A.float() @ B.float().t(), K)
    assert Cbig[:, N:].abs().max().item() == 0


def test_gemm_mn_major_unaligned_extent():
    """wgrad with an MN-major operand whose contiguous extent is not a multiple of 64 (tp-sharded vocab)."""
    from b200train.ops import gemm

    T, V, H = 512, 16032 // 8, 256   # 2004 columns: % 64 == 20
    dy, x = _mk(T, V, 11), _mk(T, H, 12)
    dw = gemm(dy, x, a_mn=True, b_mn=True)
    _check(dw, dy.float().t() @ x.float(), T)
    acc = _mk(V, H, 13)
    acc0 = acc.float().clone()
    gemm(dy, x, a_mn=True, b_mn=True, out=acc, epilogue=3)
    _check(acc, acc0 + dy.float().t() @ x.float(), T)


@pytest.mark.parametrize("config", [1 | (256 << 8), 2 | (256 << 8)])
@pytest.mark.parametrize("shape", [(4096, 4096, 256), (3840, 4160, 320)])
@pytest.mark.parametrize("layout", ["kk", "kmn", "mnmn"])
def test_gemm_tail_split(config, shape, layout):
    """Shapes whose last wave is less than half full (256 or 255 tile pairs on 74 workers, 512 / 510
    tiles on 148): the remainder runs as half-width units. Same values as the whole-tile schedule."""
    from b200train.ops import gemm

    M, N, K = shape
    a_mn = layout == "mnmn"
    b_mn = layout in ("kmn", "mnmn")
    A, B = _mk(M, K, 11), _mk(N, K, 12)
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    split = gemm(a, b, a_mn=a_mn, b_mn=b_mn, config=config)
    whole = gemm(a, b, a_mn=a_mn, b_mn=b_mn, config=config | (1 << 24))
    torch.cuda.synchronize()
    _check(split, A.float() @ B.float().t(), K)
    assert torch.equal(split, whole)
    # accumulate epilogue through the split tail
    C = _mk(M, N, 13)
    C0 = C.float().clone()
    gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=C, epilogue=3, config=config)
    _check(C, A.float() @ B.float().t() + C0, K)
