"""Differentiable ops backed by the sm_100a kernels in ``csrc/`` (PyTorch reference math on CPU)."""
from .activation import gelu_new, swiglu
from .attention import attention
from .cross_entropy import (IGNORE, linear_cross_entropy, shift_labels, vocab_parallel_cross_entropy)
from .dropout import dropout
from .embedding import embedding
from .fused import linear_rope, mlp_swiglu
from .linear import gemm, linear
from .norm import layernorm, rmsnorm
from .optim import adamw_step
from .rope import rope_qk, rope_tables

__all__ = [
    "gelu_new", "swiglu", "attention", "IGNORE", "linear_cross_entropy", "shift_labels",
    "vocab_parallel_cross_entropy", "embedding", "gemm", "linear", "layernorm", "rmsnorm", "adamw_step",
    "rope_qk", "rope_tables", "linear_rope", "mlp_swiglu", "dropout",
]
