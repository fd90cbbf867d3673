"""ctypes binding of ``libb200train.so`` (C ABI declared in ``include/b200train.h``).

The library is loaded lazily, on the first call that needs a kernel. On a machine with a GPU a missing
or unloadable library is a hard error (no silent fallback to PyTorch ops); on a CPU-only machine the
ops modules never reach this file because they dispatch CPU tensors to their PyTorch reference math.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_void_p
from pathlib import Path

import torch

_LIB = None
_LAUNCHES = 0
# kernels launched per library call (for the bench's launch count)
_LAUNCH_MULT = {"b2_rmsnorm_bwd": 2, "b2_layernorm_bwd": 2, "b2_attn_bwd": 3, "b2_attn_bwd_dropout": 3, "b2_rmsnorm_bwd_partial_rows": 0,
                "b2_colsum": 2, "b2_colsum_partial_rows": 0}
_LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libb200train.so"


class KernelError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def _declare(lib):
    P, I, F, Z = c_void_p, c_int, c_float, c_size_t
    sig = {
        "b2_gemm_bf16": [P, P, P, I, I, I, I, I, I, I, I, I, P, I, I, P],
        "b2_attn_fwd": [P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, F, I, P],
        "b2_attn_bwd": [P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, I, I, I, F, I, P],
        "b2_rmsnorm_bwd_partial_rows": [I],
        "b2_rmsnorm_fwd": [P, P, P, P, P, P, I, I, F, P],
        "b2_rmsnorm_bwd": [P, P, P, P, P, P, P, P, I, I, I, P],
        "b2_layernorm_fwd": [P, P, P, P, P, P, P, P, I, I, F, P],
        "b2_layernorm_bwd": [P, P, P, P, P, P, P, P, P, P, I, I, I, P],
        "b2_swiglu_fwd": [P, P, I, I, P],
        "b2_swiglu_bwd": [P, P, P, I, I, P],
        "b2_gelu_fwd": [P, P, Z, P],
        "b2_gelu_bwd": [P, P, P, Z, P],
        "b2_rope": [P, P, P, P, I, I, I, I, I, I, P],
        "b2_embedding_fwd": [P, P, P, I, I, P],
        "b2_embedding_bwd": [P, P, P, P, I, I, I, P],
        "b2_ce_fwd": [P, P, P, P, I, I, I, I, P],
        "b2_ce_shard_stats": [P, P, P, P, P, I, I, I, I, I, P],
        "b2_ce_bwd": [P, P, P, P, I, I, I, I, I, F, P, P],
        "b2_adamw": [P, P, P, P, Z, F, F, F, F, F, I, F, I, P],
        "b2_adamw_multicast": [P, P, P, P, P, Z, F, F, F, F, F, I, F, I, P],
        "b2_add": [P, P, P, Z, P],
        "b2_dropout": [P, P, Z, F, ctypes.c_ulonglong, ctypes.c_ulonglong, P],
        "b2_attn_fwd_dropout": [P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, F, I, F, ctypes.c_ulonglong,
                                ctypes.c_ulonglong, P],
        "b2_attn_bwd_dropout": [P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, I, I, I, F, I,
                                F, ctypes.c_ulonglong, ctypes.c_ulonglong, P],
        "b2_attn_dropout_mask": [P, I, I, I, F, ctypes.c_ulonglong, ctypes.c_ulonglong, P],
        "b2_colsum_partial_rows": [I, I],
        "b2_colsum": [P, P, P, I, I, I, I, P],
        "b2_gemm_ag": [P, P, P, I, I, I, I, I, I, I, P, ctypes.c_uint, I, I, I, P],
        "b2_gemm_rs": [P, P, P, I, I, I, I, I, I, I, I, I, I, P],
        "b2_sum_slots": [P, P, Z, I, P],
        "b2_gemm_bf16_bwait": [P, P, P, I, I, I, I, I, I, I, I, P, P, ctypes.c_uint, ctypes.c_longlong,
                               ctypes.c_longlong, I, I, I, I, P],
        "b2_gemm_swiglu": [P, P, P, P, I, I, I, I, I, I, I, P, ctypes.c_uint, ctypes.c_longlong, ctypes.c_longlong,
                           I, I, I, I, P],
        "b2_gemm_dswiglu": [P, P, P, P, I, I, I, I, I, I, I, P, ctypes.c_uint, ctypes.c_longlong, ctypes.c_longlong,
                            I, I, I, I, P],
        "b2_gemm_rope": [P, P, P, I, I, I, I, I, I, P, P, P, I, I, P, ctypes.c_uint, ctypes.c_longlong,
                         ctypes.c_longlong, I, I, I, I, P],
        "b2_peer_all_gather_flagged": [P, P, Z, I, I, I, I, I, P, ctypes.c_uint, P],
        "b2_stream_wait_flags": [P, I, I, ctypes.c_uint, P],
        "b2_local_copy": [P, P, Z, I, P],
        "b2_peer_all_gather_chunks": [P, P, Z, I, I, P, ctypes.c_uint, P],
        "b2_cpu_adamw_bf16": [P, P, P, P, Z, F, F, F, F, F, I, F],
        "b2_cpu_set_threads": [I],
        "b2_cpu_dropout": [P, P, Z, I, F, ctypes.c_ulonglong, ctypes.c_ulonglong],
        "b2_peer_barrier": [P, I, I, I, ctypes.c_uint, P],
        "b2_peer_all_gather": [P, P, Z, I, I, I, I, P],
        "b2_peer_all_gather_ce": [P, P, Z, I, I, I, I, P],
        "b2_comm_set_timeout_ms": [ctypes.c_ulonglong],
        "b2_ce_combine_peer": [P, P, P, P, I, I, I, P],
        "b2_peer_pull_2d": [P, Z, Z, P, Z, Z, Z, Z, I, I, P],
        "b2_signal_peers": [P, I, I, I, ctypes.c_uint, P],
        "b2_stream_wait_pad": [P, I, I, I, ctypes.c_uint, P],
        "b2_ce_pull_slices": [P, P, Z, I, I, P],
        "b2_reduce_slices": [P, P, P, I, Z, F, I, P],
        "b2_reduce_adamw": [P, P, P, P, P, I, Z, F, F, F, F, F, F, I, I, P],
        "b2_peer_reduce_scatter": [P, P, Z, I, I, F, P, I, ctypes.c_uint, P, I, P],
        "b2_peer_reduce_scatter_adamw": [P, P, P, P, Z, I, I, F, F, F, F, F, F, I, P, I, ctypes.c_uint, P, I, P],
        "b2_peer_all_reduce": [P, Z, I, I, F, P, I, ctypes.c_uint, P, I, P],
        "b2_nvls_reduce_scatter": [P, P, Z, I, I, F, P, I, ctypes.c_uint, P, I, P],
        "b2_nvls_reduce_scatter_adamw": [P, P, P, P, Z, I, I, F, F, F, F, F, F, I, P, I, ctypes.c_uint, P, I, P],
        "b2_nvls_all_reduce": [P, Z, I, I, F, P, I, ctypes.c_uint, P, I, P],
    }
    for name, argtypes in sig.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue  # optional symbol (checked at call time)
        fn.argtypes = argtypes
        fn.restype = c_int
    lib.b2_last_error.restype = c_char_p
    lib.b2_last_error.argtypes = []


def load():
    """Load (building first if the .so is missing and nvcc is present) and return the library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if os.environ.get("B200TRAIN_NO_AUTOBUILD") == "1":
        if not _LIB_PATH.exists():
            raise KernelError(f"{_LIB_PATH} is missing; run `python -m b200train.build`")
    else:
        # digest-incremental: a no-op when csrc/ is unchanged, a rebuild when the library is missing or stale
        # (the .so is git-ignored, so a tree can carry a library older than its sources). All ranks of a job
        # reach this at once: the build itself is serialised by a file lock (see build.build).
        from . import build as _build

        try:
            _build.build()
        except RuntimeError:
            if not _LIB_PATH.exists():
                raise
    try:
        lib = ctypes.CDLL(str(_LIB_PATH))
    except OSError as e:  # pragma: no cover - depends on the machine
        raise KernelError(f"cannot load {_LIB_PATH}: {e}") from e
    _declare(lib)
    _LIB = lib
    return lib


def available() -> bool:
    """True when kernels can run here: a CUDA device is visible and the library loads."""
    if not torch.cuda.is_available():
        return False
    load()
    return True


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t):
    """Device pointer of a tensor (or None)."""
    if t is None:
        return None
    return t.data_ptr()


def call(name: str, *args):
    """Invoke ``name`` from the library; raise KernelError with the library's message on failure."""
    lib = load()
    fn = getattr(lib, name, None)
    if fn is None:
        raise KernelError(f"{name} is not exported by {_LIB_PATH}")
    global _LAUNCHES
    _LAUNCHES += _LAUNCH_MULT.get(name, 1)
    rc = fn(*args)
    if rc != 0:
        msg = lib.b2_last_error()
        raise KernelError(f"{name} failed (rc={rc}): {msg.decode() if msg else 'unknown error'}")
    return rc


def launch_count() -> int:
    """Number of b200train kernels launched so far by this process."""
    return _LAUNCHES
