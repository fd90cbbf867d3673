"""Llama-family decoder built from the b200train ops.

Differences from the HF module tree the reference instantiates (``AutoModelForCausalLM.from_config``,
``01-single-gpu/train_llm.py:48-49``), all internal:
  * q/k/v projections are one fused [Hq*D + 2*Hkv*D, H] weight and gate/up one fused [2I, H] weight, so
    each is a single tcgen05 GEMM; ``state_dict()`` still exposes the HF names
    (``model.layers.N.self_attn.q_proj.weight`` ...) so checkpoints, TP plans and wrap policies line up.
  * activations are kept flattened [T = B*S, H]; sequences are described by ``cu_seqlens`` so padded-free
    packed batches (``00-rime/packed_dataset.py``) and fixed-length batches share one path.
  * the loss is the fused head + cross-entropy op (no fp32 [T, V] logits).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn

from .. import ops
from .config import ModelConfig


@dataclass
class CausalLMOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None


class FusedLinearWeight(nn.Module):
    """Holds a [out, in] weight; ``parts`` names the HF sub-matrices stacked along ``out``."""

    def __init__(self, in_features, out_features, dtype, device, parts=None):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_features, in_features, dtype=dtype, device=device))
        self.weight._b2_linear_only = True   # read by ops.linear GEMMs only (FSDP: gather fused into them)
        self.parts = parts or []

    def forward(self, x):
        return ops.linear(x, self.weight)


class RMSNorm(nn.Module):
    def __init__(self, hidden, eps, dtype, device):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        return ops.rmsnorm(x, self.weight, self.eps, residual)

    def reset_parameters(self):  # reference shim: 04-fully-sharded-data-parallel/train_llm.py:40-44
        nn.init.ones_(self.weight)


class LlamaAttention(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        self.cfg = cfg
        self.n_q, self.n_kv, self.D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        self.qkv_proj = FusedLinearWeight(cfg.hidden_size, cfg.q_size + 2 * cfg.kv_size, dtype, device,
                                          parts=[("q_proj", cfg.q_size), ("k_proj", cfg.kv_size),
                                                 ("v_proj", cfg.kv_size)])
        self.o_proj = FusedLinearWeight(cfg.q_size, cfg.hidden_size, dtype, device)

    def forward(self, x, ctx):
        # QKV projection with the rotary embedding of the q / k heads in the GEMM epilogue
        qkv = ops.linear_rope(x, self.qkv_proj.weight, ctx.cos, ctx.sin, ctx.pos, self.n_q + self.n_kv, self.D,
                              ctx.seq_len)
        o = ops.attention(qkv, ctx.cu_seqlens, ctx.max_seqlen, self.n_q, self.n_kv, self.D)
        return self.o_proj(o)


class LlamaMLP(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        I = cfg.intermediate_size
        self.gate_up_proj = FusedLinearWeight(cfg.hidden_size, 2 * I, dtype, device,
                                              parts=[("gate_proj", I), ("up_proj", I)])
        self.down_proj = FusedLinearWeight(I, cfg.hidden_size, dtype, device)

    def forward(self, x):
        # gate/up GEMM with SwiGLU in its epilogue, down projection, and the mirrored fused backward
        return ops.mlp_swiglu(x, self.gate_up_proj.weight, self.down_proj.weight)


class LlamaDecoderLayer(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        self.input_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)
        self.self_attn = LlamaAttention(cfg, dtype, device)
        self.post_attention_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)
        self.mlp = LlamaMLP(cfg, dtype, device)

    def forward(self, h, delta, ctx):
        """``h``: residual stream, ``delta``: previous block output not yet added (or None)."""
        x, h = self.input_layernorm(delta if delta is not None else h, h if delta is not None else None)
        a = self.self_attn(x, ctx)
        x, h = self.post_attention_layernorm(a, h)
        return h, self.mlp(x)


@dataclass
class SeqContext:
    cos: torch.Tensor
    sin: torch.Tensor
    pos: Optional[torch.Tensor]
    seq_len: int
    cu_seqlens: torch.Tensor
    max_seqlen: int


class LlamaModel(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device)
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, dtype, device) for _ in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, dtype, device)


class LlamaForCausalLM(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.config = cfg
        self.model = LlamaModel(cfg, dtype, device)
        self.lm_head = nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False, dtype=dtype, device=device)
        if cfg.tie_word_embeddings:
            self.lm_head.weight = self.model.embed_tokens.weight
        self._rope = None
        # parallel strategies (FSDP gather/scatter, activation checkpointing) install a runner:
        # unit_runner(unit_index, fn, *tensors) -> fn(*tensors); unit 0 = embedding, 1..L = layers,
        # L+1 = final norm + head (same order as units())
        self.unit_runner = None
        self.layer_groups = None  # None: every layer is its own unit (see group_layers)
        self.layer_groups = None  # None: every layer is its own unit (see group_layers)

    # ---- init (HF _init_weights: normal(0, initializer_range) for linears / embeddings, ones for norms)
    @torch.no_grad()
    def init_weights(self, seed: Optional[int] = None):
        gen = None
        if seed is not None:
            dev = self.lm_head.weight.device
            gen = torch.Generator(device=dev if dev.type != "meta" else "cpu").manual_seed(seed)
        std = self.config.initializer_range
        for name, p in self.named_parameters():
            if p.device.type == "meta":
                continue
            if name.endswith("norm.weight") or "layernorm" in name:
                p.fill_(1.0)
            else:
                p.normal_(0.0, std, generator=gen)

    # ---- units for flat storage / FSDP wrapping (reference wrap policy: decoder layers + embedding,
    #      05-training-llama-405b/train_llm.py:101-107)
    def units(self):
        out = [("model.embed_tokens", [self.model.embed_tokens.weight])]
        for g in self._groups():
            name = f"model.layers.{g[0]}" if len(g) == 1 else f"model.layers.{g[0]}-{g[-1]}"
            out.append((name, [p for j in g for p in self.model.layers[j].parameters()]))
        tail = [self.model.norm.weight]
        if not self.config.tie_word_embeddings:
            tail.append(self.lm_head.weight)
        out.append(("model.norm+lm_head", tail))
        return out

    def _groups(self):
        return self.layer_groups or [[i] for i in range(len(self.model.layers))]

    def group_layers(self, min_numel: int):
        """Merge consecutive layers into wrap units of >= ``min_numel`` parameters (``--numel-to-wrap``,
        ``04-fully-sharded-data-parallel/train_llm.py:84-86``). Must be called before a strategy lays
        out its flat storage."""
        from .config import group_by_numel

        sizes = [sum(p.numel() for p in layer.parameters()) for layer in self.model.layers]
        self.layer_groups = group_by_numel(sizes, min_numel) if min_numel > 0 else None

    def _rope_tables(self, device, n):
        if self._rope is None or self._rope[0].device != device or self._rope[0].shape[0] < n:
            cfg = self.config
            n_alloc = max(n, min(cfg.max_position_embeddings, 8192))
            self._rope = ops.rope_tables(cfg.head_dim, n_alloc, cfg.rope_theta, cfg.rope_scaling, device=device)
        return self._rope

    def make_context(self, batch_size, seq_len, device, position_ids=None) -> SeqContext:
        cos, sin = self._rope_tables(device, seq_len)
        if position_ids is None:
            cu = torch.arange(0, batch_size + 1, dtype=torch.int32, device=device) * seq_len
            return SeqContext(cos, sin, None, seq_len, cu, seq_len)
        pos = position_ids.reshape(-1).to(torch.int32)
        starts = torch.nonzero(pos == 0).flatten().to(torch.int32)
        total = torch.tensor([pos.numel()], dtype=torch.int32, device=device)
        cu = torch.cat((starts, total))
        max_len = int((cu[1:] - cu[:-1]).max().item())
        return SeqContext(cos, sin, pos.contiguous(), seq_len, cu, max_len)

    def _run(self, idx, fn, *tensors):
        if self.unit_runner is None:
            return fn(*tensors)
        return self.unit_runner(idx, fn, *tensors)

    def forward(self, input_ids, labels=None, attention_mask=None, position_ids=None, **_):
        B, S = input_ids.shape
        ctx = self.make_context(B, S, input_ids.device, position_ids)
        flat_ids = input_ids.reshape(-1)
        h = self._run(0, lambda: ops.embedding(flat_ids, self.model.embed_tokens.weight))
        delta = None
        layers, groups = self.model.layers, self._groups()
        for gi, g in enumerate(groups):
            def run_group(a, b=None, g=g):
                for j in g:
                    a, b = layers[j](a, b, ctx)
                return a, b

            h, delta = self._run(gi + 1, run_group, h) if delta is None else self._run(gi + 1, run_group, h, delta)
        L = len(groups)
        shifted = ops.shift_labels(labels).reshape(-1) if labels is not None else None

        def head(a, b=None):
            x, _ = self.model.norm(b, a) if b is not None else self.model.norm(a)
            if shifted is None:
                return ops.linear(x, self.lm_head.weight)
            return ops.linear_cross_entropy(x, self.lm_head.weight, shifted)

        out = self._run(L + 1, head, h, delta) if delta is not None else self._run(L + 1, head, h)
        if labels is None:
            return CausalLMOutput(logits=out.view(B, S, -1))
        return CausalLMOutput(loss=out)

    # ---- HF-compatible state dict (fused weights are exposed as their parts; views, no copies)
    def hf_state_dict(self):
        sd = {}
        for name, mod in self.named_modules():
            if isinstance(mod, FusedLinearWeight):
                parent = name.rsplit(".", 1)[0]
                if mod.parts:
                    off = 0
                    for part, n in mod.parts:
                        sd[f"{parent}.{part}.weight"] = mod.weight.data[off:off + n]
                        off += n
                else:
                    sd[f"{name}.weight"] = mod.weight.data
        for name, p in self.named_parameters():
            if "qkv_proj" in name or "gate_up_proj" in name or "o_proj" in name or "down_proj" in name:
                continue
            if name == "lm_head.weight" and self.config.tie_word_embeddings:
                continue
            sd[name] = p.data
        if self.config.tie_word_embeddings:
            sd["lm_head.weight"] = self.model.embed_tokens.weight.data
        return sd

    @torch.no_grad()
    def load_hf_state_dict(self, sd, strict=True):
        own = self.hf_state_dict()
        missing = [k for k in own if k not in sd]
        if strict and missing:
            raise KeyError(f"missing keys: {missing[:5]} ...")
        for k, dst in own.items():
            if k in sd:
                dst.copy_(sd[k])
        return missing
