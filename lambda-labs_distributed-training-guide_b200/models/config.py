"""Model configuration: presets for the architectures the reference trains, plus loading of a
Hugging Face ``config.json`` from a local directory (there is no hub access on the training box).

Reference: ``AutoConfig.from_pretrained(args.model_name, use_cache=False)``
(``01-single-gpu/train_llm.py:47``); model ids used by the guide: ``openai-community/gpt2``,
``meta-llama/Llama-2-7b-hf``, ``meta-llama/Llama-3.1-8B/70B/405B``, ``meta-llama/Llama-3.2-3B-Instruct``.
"""
from __future__ import annotations

import json
from dataclasses import asdict, dataclass
from pathlib import Path
from typing import Optional


@dataclass
class ModelConfig:
    model_type: str = "llama"            # "llama" | "gpt2"
    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: Optional[int] = None
    max_position_embeddings: int = 131072
    rms_norm_eps: float = 1e-5
    layer_norm_epsilon: float = 1e-5
    rope_theta: float = 500000.0
    rope_scaling: Optional[dict] = None
    tie_word_embeddings: bool = False
    initializer_range: float = 0.02
    # GPT-2 dropout, one rate for HF's embd_pdrop / attn_pdrop / resid_pdrop (all 0.1 in GPT2Config's defaults)
    dropout: float = 0.0
    name: str = ""

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_attention_heads

    # ---- derived sizes ----
    @property
    def q_size(self):
        return self.num_attention_heads * self.head_dim

    @property
    def kv_size(self):
        return self.num_key_value_heads * self.head_dim

    def num_parameters(self) -> int:
        H, I, V, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        if self.model_type == "gpt2":
            per_layer = (H * 3 * H + 3 * H) + (H * H + H) + (H * I + I) + (I * H + H) + 4 * H
            return V * H + self.max_position_embeddings * H + L * per_layer + 2 * H
        per_layer = H * (self.q_size + 2 * self.kv_size) + self.q_size * H + 3 * H * I + 2 * H
        head = 0 if self.tie_word_embeddings else V * H
        return V * H + L * per_layer + H + head

    def flops_per_token(self, seq_len: int) -> float:
        """6 * N_non_embedding + 12 * L * H * S (BASELINE.md definition, full S x S attention)."""
        H, V, L = self.hidden_size, self.vocab_size, self.num_hidden_layers
        n = self.num_parameters() - V * H
        if self.model_type == "gpt2":
            n -= self.max_position_embeddings * H
            n += V * H  # tied head matmul still costs FLOPs
        return 6.0 * n + 12.0 * L * H * seq_len

    def to_dict(self):
        return asdict(self)


_LLAMA3_SCALING = {"factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                   "original_max_position_embeddings": 8192, "rope_type": "llama3"}

PRESETS = {
    "gpt2": dict(model_type="gpt2", vocab_size=50257, hidden_size=768, intermediate_size=3072,
                 num_hidden_layers=12, num_attention_heads=12, num_key_value_heads=12,
                 max_position_embeddings=1024, tie_word_embeddings=True, dropout=0.1),
    "llama-2-7b": dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                       num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096,
                       rope_theta=10000.0),
    "llama-3.2-3b": dict(vocab_size=128256, hidden_size=3072, intermediate_size=8192, num_hidden_layers=28,
                         num_attention_heads=24, num_key_value_heads=8, head_dim=128,
                         rope_scaling=dict(_LLAMA3_SCALING, factor=32.0), tie_word_embeddings=True),
    "llama-3.1-8b": dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                         num_attention_heads=32, num_key_value_heads=8, rope_scaling=_LLAMA3_SCALING),
    "llama-3.1-70b": dict(vocab_size=128256, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                          num_attention_heads=64, num_key_value_heads=8, rope_scaling=_LLAMA3_SCALING),
    "llama-3.1-405b": dict(vocab_size=128256, hidden_size=16384, intermediate_size=53248,
                           num_hidden_layers=126, num_attention_heads=128, num_key_value_heads=8,
                           rope_scaling=_LLAMA3_SCALING),
    # tiny shapes for tests / smoke (same code paths, head_dim 64 and 128)
    "llama-tiny": dict(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=2, head_dim=64, max_position_embeddings=2048,
                       rope_theta=10000.0),
    "llama-smoke": dict(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                        num_attention_heads=4, num_key_value_heads=2, head_dim=128, max_position_embeddings=4096),
    "gpt2-tiny": dict(model_type="gpt2", vocab_size=512, hidden_size=128, intermediate_size=512,
                      num_hidden_layers=2, num_attention_heads=2, num_key_value_heads=2,
                      max_position_embeddings=256, tie_word_embeddings=True),
}

_ALIASES = {
    "openai-community/gpt2": "gpt2",
    "meta-llama/llama-2-7b-hf": "llama-2-7b",
    "meta-llama/llama-3.1-8b": "llama-3.1-8b",
    "meta-llama/meta-llama-3.1-8b": "llama-3.1-8b",
    "meta-llama/llama-3.1-70b": "llama-3.1-70b",
    "meta-llama/meta-llama-3.1-70b": "llama-3.1-70b",
    "meta-llama/llama-3.1-405b": "llama-3.1-405b",
    "meta-llama/meta-llama-3.1-405b": "llama-3.1-405b",
    "meta-llama/llama-3.2-3b-instruct": "llama-3.2-3b",
    "meta-llama/llama-3.2-3b": "llama-3.2-3b",
}


def _from_hf_dict(d: dict, name: str) -> ModelConfig:
    mt = d.get("model_type", "llama")
    if mt == "gpt2":
        H = d.get("n_embd", 768)
        return ModelConfig(model_type="gpt2", vocab_size=d.get("vocab_size", 50257), hidden_size=H,
                           intermediate_size=d.get("n_inner") or 4 * H,
                           num_hidden_layers=d.get("n_layer", 12), num_attention_heads=d.get("n_head", 12),
                           num_key_value_heads=d.get("n_head", 12),
                           max_position_embeddings=d.get("n_positions", 1024),
                           layer_norm_epsilon=d.get("layer_norm_epsilon", 1e-5),
                           tie_word_embeddings=d.get("tie_word_embeddings", True),
                           dropout=float(d.get("resid_pdrop", 0.1)),   # GPT2Config default (one rate for the 3 sites)
                           initializer_range=d.get("initializer_range", 0.02), name=name)
    rope_scaling = d.get("rope_scaling")
    rope_theta = d.get("rope_theta", 10000.0)
    rp = d.get("rope_parameters")
    if rp:  # transformers >= 5 layout
        rope_theta = rp.get("rope_theta", rope_theta)
        if rp.get("rope_type", "default") != "default":
            rope_scaling = rp
    return ModelConfig(model_type="llama", vocab_size=d["vocab_size"], hidden_size=d["hidden_size"],
                       intermediate_size=d["intermediate_size"], num_hidden_layers=d["num_hidden_layers"],
                       num_attention_heads=d["num_attention_heads"],
                       num_key_value_heads=d.get("num_key_value_heads", d["num_attention_heads"]),
                       head_dim=d.get("head_dim"), max_position_embeddings=d.get("max_position_embeddings", 4096),
                       rms_norm_eps=d.get("rms_norm_eps", 1e-5), rope_theta=rope_theta, rope_scaling=rope_scaling,
                       tie_word_embeddings=d.get("tie_word_embeddings", False),
                       initializer_range=d.get("initializer_range", 0.02), name=name)


def load_config(name_or_path: str, **overrides) -> ModelConfig:
    """Resolve ``--model-name``: a local directory with ``config.json``, a preset, or a known hub id."""
    p = Path(name_or_path)
    if p.is_dir() and (p / "config.json").exists():
        cfg = _from_hf_dict(json.loads((p / "config.json").read_text()), name_or_path)
    elif p.is_file() and p.suffix == ".json":
        cfg = _from_hf_dict(json.loads(p.read_text()), name_or_path)
    else:
        key = name_or_path.lower()
        key = _ALIASES.get(key, key)
        if key not in PRESETS:
            raise ValueError(
                f"unknown model '{name_or_path}': pass a directory containing config.json or one of "
                f"{sorted(PRESETS)} (hub ids {sorted(_ALIASES)} map to presets; the box is offline)")
        cfg = ModelConfig(**PRESETS[key], name=key)
    for k, v in overrides.items():
        if v is not None:
            setattr(cfg, k, v)
    cfg.__post_init__()
    return cfg


def to_hf_dict(cfg: ModelConfig) -> dict:
    """A ``config.json`` payload that HF ``AutoConfig`` accepts (used to drive the reference arm)."""
    if cfg.model_type == "gpt2":
        return dict(model_type="gpt2", architectures=["GPT2LMHeadModel"], vocab_size=cfg.vocab_size,
                    n_embd=cfg.hidden_size, n_layer=cfg.num_hidden_layers, n_head=cfg.num_attention_heads,
                    n_positions=cfg.max_position_embeddings, n_inner=cfg.intermediate_size,
                    activation_function="gelu_new", layer_norm_epsilon=cfg.layer_norm_epsilon,
                    resid_pdrop=cfg.dropout, embd_pdrop=cfg.dropout, attn_pdrop=cfg.dropout,
                    initializer_range=cfg.initializer_range, bos_token_id=0, eos_token_id=0)
    d = dict(model_type="llama", architectures=["LlamaForCausalLM"], vocab_size=cfg.vocab_size,
             hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
             num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
             num_key_value_heads=cfg.num_key_value_heads, head_dim=cfg.head_dim,
             max_position_embeddings=cfg.max_position_embeddings, rms_norm_eps=cfg.rms_norm_eps,
             rope_theta=cfg.rope_theta, tie_word_embeddings=cfg.tie_word_embeddings,
             hidden_act="silu", attention_bias=False, mlp_bias=False, attention_dropout=0.0,
             initializer_range=cfg.initializer_range, bos_token_id=1, eos_token_id=2,
             torch_dtype="bfloat16")
    if cfg.rope_scaling:
        d["rope_scaling"] = dict(cfg.rope_scaling)
    return d


def group_by_numel(sizes, min_numel: int):
    """Greedy grouping of consecutive layers into wrap units of at least ``min_numel`` parameters.

    Stands in for ``size_based_auto_wrap_policy(min_num_params=--numel-to-wrap)``
    (``04-fully-sharded-data-parallel/train_llm.py:84-86``): a layer that is large enough is its own unit,
    smaller layers are merged with their successors (GPT-2's twelve 7 M-parameter blocks become one unit at
    the default 100 M, as they end up in the root wrap of the reference). A short tail joins the last group.
    """
    groups, cur, acc = [], [], 0
    for i, n in enumerate(sizes):
        cur.append(i)
        acc += n
        if acc >= min_numel:
            groups.append(cur)
            cur, acc = [], 0
    if cur:
        if groups and min_numel > 0:
            groups[-1].extend(cur)
        else:
            groups.append(cur)
    return groups
