"""RMSNorm backward without caller-provided scratch (``b2_rmsnorm_bwd_static``).

Some users of this library cannot allocate device memory at the point where they need the kernel: a
plain-C plugin inside an application that has given all remaining device memory to its own pool, or a
latency-critical section that cannot afford cudaMalloc's device-wide synchronisation. For them the
library carries its own fixed 32 MiB scratch array. This test drives that entry point through the raw C
ABI (ctypes, no b200train op wrappers) and checks that it produces what the ordinary entry point
produces with a caller-allocated scratch buffer -- twice in a row with different inputs, because all
calls share the one array and the second call starts from the first call's leftovers.

The file name sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import ctypes
import gc
from ctypes import c_float, c_int, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from b200train import _C

    _C.load()   # builds the library first when it is missing or older than csrc/ (fresh checkout, stale .so)
    lib = ctypes.CDLL(str(_C.lib_path()))
    P = c_void_p
    lib.b2_rmsnorm_bwd_static.argtypes = [P, P, P, P, P, P, P, c_int, c_int, c_int, P]
    lib.b2_rmsnorm_bwd_static.restype = c_int
    lib.b2_rmsnorm_bwd.argtypes = [P, P, P, P, P, P, P, P, c_int, c_int, c_int, P]
    lib.b2_rmsnorm_bwd.restype = c_int
    lib.b2_rmsnorm_fwd.argtypes = [P, P, P, P, P, P, c_int, c_int, c_float, P]
    lib.b2_rmsnorm_fwd.restype = c_int
    lib.b2_rmsnorm_bwd_partial_rows.argtypes = [c_int]
    lib.b2_rmsnorm_bwd_partial_rows.restype = c_int
    lib.b2_rmsnorm_bwd_static_capacity.argtypes = []
    lib.b2_rmsnorm_bwd_static_capacity.restype = c_size_t
    lib.b2_last_error.restype = ctypes.c_char_p
    return lib


def _inputs(T, H, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda *s: torch.randn(*s, device="cuda", generator=g).to(torch.bfloat16)  # noqa: E731
    return mk(T, H), mk(T, H), (1 + 0.1 * mk(H).float()).to(torch.bfloat16), mk(T, H)


def _close(a, b, atol, rtol=2e-2):
    a, b = a.float(), b.float()
    assert bool(((a - b).abs() <= atol + rtol * b.abs()).all()), (a - b).abs().max().item()


def test_static_scratch_matches_caller_scratch_two_calls_in_a_row():
    lib = _lib()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    T, H = 1500, 4096
    nb = lib.b2_rmsnorm_bwd_partial_rows(T)
    assert nb * H <= lib.b2_rmsnorm_bwd_static_capacity()
    for seed in (11, 12):  # second pass: different data over the first pass's leftovers
        h, dy, w, dres = _inputs(T, H, seed)
        y = torch.empty_like(h)
        rstd = torch.empty(T, dtype=torch.float32, device="cuda")
        assert lib.b2_rmsnorm_fwd(y.data_ptr(), rstd.data_ptr(), None, h.data_ptr(), None, w.data_ptr(), T, H,
                                  1e-5, stream) == 0
        # existing entry point, scratch allocated by the caller (here: by torch)
        partial = torch.empty(nb * H, dtype=torch.float32, device="cuda")
        dx_ref, dw_ref = torch.empty_like(h), torch.empty_like(w)
        rc = lib.b2_rmsnorm_bwd(dx_ref.data_ptr(), dw_ref.data_ptr(), partial.data_ptr(), dy.data_ptr(),
                                h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(), T, H, 0, stream)
        assert rc == 0, lib.b2_last_error()
        # new entry point: outputs are allocated up front, the call itself allocates nothing
        dx, dw = torch.empty_like(h), torch.empty_like(w)
        gc.collect()                       # leftovers of earlier tests must not be freed mid-measurement
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]     # device-level free bytes (sees any cudaMalloc)
        rc = lib.b2_rmsnorm_bwd_static(dx.data_ptr(), dw.data_ptr(), dy.data_ptr(), h.data_ptr(), w.data_ptr(),
                                       rstd.data_ptr(), dres.data_ptr(), T, H, 0, stream)
        assert rc == 0, lib.b2_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before, "the no-allocation entry point allocated device memory"
        _close(dx, dx_ref, 3e-2)
        _close(dw, dw_ref, 0.3, 3e-2)


def test_static_scratch_rejects_oversize():
    lib = _lib()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    T, H = 4096, 16384  # partial rows x H exceeds the built-in 8 Mi floats
    assert lib.b2_rmsnorm_bwd_partial_rows(T) * H > lib.b2_rmsnorm_bwd_static_capacity()
    buf = torch.zeros(8, device="cuda")
    rc = lib.b2_rmsnorm_bwd_static(buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr(),
                                   buf.data_ptr(), None, T, H, 0, stream)
    assert rc != 0
    assert b"built-in" in lib.b2_last_error()
