"""b200train: a Blackwell (sm_100a) native causal-LM training framework.

Capabilities follow the chapter scripts of rimelabs/lambda-labs_distributed-training-guide
(single GPU, DDP + ZeRO-1, FSDP full-shard with CPU offload and activation checkpointing, tensor
parallel with sequence / loss parallel, 2D FSDP x TP), with the hot ops implemented as hand-written
CUDA kernels in ``csrc/`` (tcgen05 GEMM and attention, fused norms / RoPE / SwiGLU / CE / AdamW and
NVLink peer-memory collectives).
"""
__version__ = "0.1.0"
