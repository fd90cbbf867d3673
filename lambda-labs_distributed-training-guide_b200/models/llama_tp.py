"""Tensor-parallel Llama with sequence parallel and loss parallel.

The plan of the reference (``06-tensor-parallel/train_llm.py:84-128``), expressed on local shards instead
of DTensors:
  embed_tokens      hidden-sharded table [V, H/tp]; lookup result exchanged to sequence-sharded [T/tp, H]
  input/post norms  sequence parallel: run on the local token rows, weights replicated
  q/k/v, gate/up    column parallel: local heads / local intermediate columns; input all-gathered on the
                    sequence in front of the GEMM
  o_proj, down      row parallel: partial sums reduce-scattered on the sequence
  lm_head           column parallel over the vocabulary; logits stay vocabulary-sharded and feed the
                    vocabulary-parallel cross entropy (``loss_parallel``, ``:239-243``)
Attention runs on ``n_heads/tp`` local heads over the full sequence, so ``n_kv_heads % tp == 0``
(``06-tensor-parallel/README.md:108``). Position ids are always explicit / derived from the full sequence
(the reference needs ``position_ids = arange(seq_length)`` for the same reason, ``:231-233``).
"""
from __future__ import annotations

import copy

import torch
from torch import nn

from .. import ops
from ..parallel.tp_fused import colwise, rowwise
from ..parallel.tp_ops import TPGroup, hidden_to_seq
from .config import ModelConfig
from .llama import CausalLMOutput, LlamaForCausalLM


def local_config(cfg: ModelConfig, tp: int) -> ModelConfig:
    assert cfg.model_type == "llama", "tensor parallel is implemented for the Llama family"
    assert not cfg.tie_word_embeddings, "tensor parallel needs untied embeddings / lm_head"
    for name, v in (("num_attention_heads", cfg.num_attention_heads), ("num_key_value_heads", cfg.num_key_value_heads),
                    ("intermediate_size", cfg.intermediate_size), ("hidden_size", cfg.hidden_size),
                    ("vocab_size", cfg.vocab_size)):
        assert v % tp == 0, f"{name}={v} is not divisible by tp={tp}"
    loc = copy.deepcopy(cfg)
    loc.num_attention_heads = cfg.num_attention_heads // tp
    loc.num_key_value_heads = cfg.num_key_value_heads // tp
    loc.intermediate_size = cfg.intermediate_size // tp
    loc.head_dim = cfg.head_dim
    return loc


class TPLlamaForCausalLM(LlamaForCausalLM):
    def __init__(self, cfg: ModelConfig, tp: TPGroup, dtype=torch.bfloat16, device=None):
        self.tp = tp
        self.global_config = cfg
        loc = local_config(cfg, tp.size)
        super().__init__(loc, dtype=dtype, device=device)
        H, V = cfg.hidden_size, cfg.vocab_size
        self.model.embed_tokens = nn.Embedding(V, H // tp.size, dtype=dtype, device=device)
        self.lm_head = nn.Linear(H, V // tp.size, bias=False, dtype=dtype, device=device)
        self.vocab_start = tp.rank * (V // tp.size)
        for name, p in self.named_parameters():
            p._b2_tp_replicated = "norm" in name  # partial gradients, summed over tp after backward

    # HF-named *global* tensors are sharded over tp along this dim (None = replicated)
    @staticmethod
    def tp_shard_dim(name: str):
        if "norm" in name:
            return None
        if name.endswith("o_proj.weight") or name.endswith("down_proj.weight") or "embed_tokens" in name:
            return 1
        return 0

    def _tp_layer(self, layer, h, delta, ctx):
        tp = self.tp
        att, mlp = layer.self_attn, layer.mlp
        x, h = layer.input_layernorm(delta if delta is not None else h, h if delta is not None else None)
        qkv = colwise(x, att.qkv_proj.weight, tp)                 # all-gather(seq) -> GEMM, local heads
        qkv = ops.rope_qk(qkv, ctx.cos, ctx.sin, ctx.pos, att.n_q + att.n_kv, att.D, ctx.seq_len)
        o = ops.attention(qkv, ctx.cu_seqlens, ctx.max_seqlen, att.n_q, att.n_kv, att.D)
        a = rowwise(o, att.o_proj.weight, tp)                     # GEMM -> reduce-scatter(seq): [T/tp, H]
        x, h = layer.post_attention_layernorm(a, h)
        gu = colwise(x, mlp.gate_up_proj.weight, tp)
        return h, rowwise(ops.swiglu(gu), mlp.down_proj.weight, tp)

    def forward(self, input_ids, labels=None, attention_mask=None, position_ids=None, **_):
        B, S = input_ids.shape
        tp = self.tp
        assert (B * S) % tp.size == 0, "tokens per step must be divisible by tp"
        ctx = self.make_context(B, S, input_ids.device, position_ids)
        flat_ids = input_ids.reshape(-1)
        h = self._run(0, lambda: hidden_to_seq(ops.embedding(flat_ids, self.model.embed_tokens.weight), tp))
        delta = None
        for i, layer in enumerate(self.model.layers):
            if delta is None:
                h, delta = self._run(i + 1, lambda a, layer=layer: self._tp_layer(layer, a, None, ctx), h)
            else:
                h, delta = self._run(i + 1, lambda a, b, layer=layer: self._tp_layer(layer, a, b, ctx), h, delta)
        L = len(self.model.layers)
        shifted = ops.shift_labels(labels).reshape(-1) if labels is not None else None

        def head(a, b=None):
            x, _ = self.model.norm(b, a) if b is not None else self.model.norm(a)
            logits = colwise(x, self.lm_head.weight, tp)                        # [T, V/tp]
            if shifted is None:
                return logits
            return ops.vocab_parallel_cross_entropy(logits, shifted, self.vocab_start, tp.group,
                                                    peer=getattr(tp, "fused", None))

        out = self._run(L + 1, head, h, delta) if delta is not None else self._run(L + 1, head, h)
        if labels is None:
            return CausalLMOutput(logits=out.view(B, S, -1))
        return CausalLMOutput(loss=out)

    @torch.no_grad()
    def load_from_full(self, full: LlamaForCausalLM):
        """Copy this rank's slices out of an unsharded model (tests / small models: identical numerics to
        the single-device model)."""
        tp, r = self.tp.size, self.tp.rank
        src = full.hf_state_dict()
        for name, dst in self.hf_state_dict().items():
            t = src[name]
            d = self.tp_shard_dim(name)
            if d is None:
                dst.copy_(t)
            else:
                n = t.shape[d] // tp
                dst.copy_(t.narrow(d, r * n, n))
