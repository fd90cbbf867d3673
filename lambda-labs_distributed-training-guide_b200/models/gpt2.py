"""GPT-2 built from the b200train ops (chapter 01/02 model of the reference:
``python train_llm.py -m openai-community/gpt2``, ``01-single-gpu/README.md:9-15``).

HF parameter names are kept (``transformer.h.N.attn.c_attn.weight`` ..., Conv1D weights stored [in, out],
``lm_head`` tied to ``wte``). HF's GPT-2 config defaults to dropout 0.1 at three sites and ``from_config``
leaves the module in training mode (``02-distributed-data-parallel/train_llm.py:60-62``): ``cfg.dropout`` is
applied to the embedding sum (embd_pdrop), to the attention probabilities inside the attention kernels
(attn_pdrop) and to both residual branches (resid_pdrop), with Philox masks regenerated in backward.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import ops
from .config import ModelConfig
from .llama import CausalLMOutput, SeqContext


class Conv1D(nn.Module):
    """HF Conv1D: y = x W + b with W [in, out]."""

    def __init__(self, nf, nx, dtype, device):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(nx, nf, dtype=dtype, device=device))
        self.weight._b2_linear_only = True   # read by ops.linear GEMMs only (FSDP: gather fused into them)
        self.bias = nn.Parameter(torch.zeros(nf, dtype=dtype, device=device))

    def forward(self, x):
        return ops.linear(x, self.weight, self.bias, w_is_kn=True)


class LayerNorm(nn.Module):
    def __init__(self, hidden, eps, dtype, device):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden, dtype=dtype, device=device))
        self.bias = nn.Parameter(torch.zeros(hidden, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        return ops.layernorm(x, self.weight, self.bias, self.eps, residual)


class GPT2Attention(nn.Module):
    def __init__(self, cfg, dtype, device):
        super().__init__()
        H = cfg.hidden_size
        self.n_head, self.D = cfg.num_attention_heads, cfg.head_dim
        self.c_attn = Conv1D(3 * H, H, dtype, device)
        self.c_proj = Conv1D(H, H, dtype, device)
        self.pdrop = cfg.dropout

    def forward(self, x, ctx):
        qkv = self.c_attn(x)
        p = self.pdrop if self.training else 0.0
        o = ops.attention(qkv, ctx.cu_seqlens, ctx.max_seqlen, self.n_head, self.n_head, self.D, dropout_p=p)
        return ops.dropout(self.c_proj(o), self.pdrop, self.training)      # attn_pdrop inside, resid_pdrop after


class GPT2MLP(nn.Module):
    def __init__(self, cfg, dtype, device):
        super().__init__()
        self.c_fc = Conv1D(cfg.intermediate_size, cfg.hidden_size, dtype, device)
        self.c_proj = Conv1D(cfg.hidden_size, cfg.intermediate_size, dtype, device)
        self.pdrop = cfg.dropout

    def forward(self, x):
        return ops.dropout(self.c_proj(ops.gelu_new(self.c_fc(x))), self.pdrop, self.training)   # resid_pdrop


class GPT2Block(nn.Module):
    def __init__(self, cfg, dtype, device):
        super().__init__()
        self.ln_1 = LayerNorm(cfg.hidden_size, cfg.layer_norm_epsilon, dtype, device)
        self.attn = GPT2Attention(cfg, dtype, device)
        self.ln_2 = LayerNorm(cfg.hidden_size, cfg.layer_norm_epsilon, dtype, device)
        self.mlp = GPT2MLP(cfg, dtype, device)

    def forward(self, h, delta, ctx):
        x, h = self.ln_1(delta if delta is not None else h, h if delta is not None else None)
        a = self.attn(x, ctx)
        x, h = self.ln_2(a, h)
        return h, self.mlp(x)


class GPT2Model(nn.Module):
    def __init__(self, cfg, dtype, device):
        super().__init__()
        self.wte = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device)
        self.wpe = nn.Embedding(cfg.max_position_embeddings, cfg.hidden_size, dtype=dtype, device=device)
        self.h = nn.ModuleList([GPT2Block(cfg, dtype, device) for _ in range(cfg.num_hidden_layers)])
        self.ln_f = LayerNorm(cfg.hidden_size, cfg.layer_norm_epsilon, dtype, device)


class GPT2LMHeadModel(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype=torch.bfloat16, device=None):
        super().__init__()
        self.config = cfg
        self.transformer = GPT2Model(cfg, dtype, device)
        self.lm_head = nn.Linear(cfg.hidden_size, cfg.vocab_size, bias=False, dtype=dtype, device=device)
        self.lm_head.weight = self.transformer.wte.weight  # tied
        self.unit_runner = None  # see LlamaForCausalLM
        self.layer_groups = None

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
            if name.endswith("bias"):
                p.zero_()
            elif ".ln_" in name or "ln_f" in name:
                p.fill_(1.0)
            else:
                p.normal_(0.0, std, generator=gen)

    def units(self):
        t = self.transformer
        out = [("transformer.wte+wpe", [t.wte.weight, t.wpe.weight])]
        for g in self._groups():
            name = f"transformer.h.{g[0]}" if len(g) == 1 else f"transformer.h.{g[0]}-{g[-1]}"
            out.append((name, [p for j in g for p in t.h[j].parameters()]))
        out.append(("transformer.ln_f", list(t.ln_f.parameters())))
        return out

    def _groups(self):
        return self.layer_groups or [[i] for i in range(len(self.transformer.h))]

    def group_layers(self, min_numel: int):
        """Merge consecutive blocks into wrap units of >= ``min_numel`` parameters (``--numel-to-wrap``)."""
        from .config import group_by_numel

        sizes = [sum(p.numel() for p in blk.parameters()) for blk in self.transformer.h]
        self.layer_groups = group_by_numel(sizes, min_numel) if min_numel > 0 else None

    @property
    def layers(self):
        return self.transformer.h

    def make_context(self, batch_size, seq_len, device, position_ids=None) -> SeqContext:
        cu = torch.arange(0, batch_size + 1, dtype=torch.int32, device=device) * seq_len
        return SeqContext(None, None, None, seq_len, cu, seq_len)

    def _run(self, idx, fn, *tensors):
        if self.unit_runner is None:
            return fn(*tensors)
        return self.unit_runner(idx, fn, *tensors)

    def forward(self, input_ids, labels=None, attention_mask=None, position_ids=None, **_):
        B, S = input_ids.shape
        t = self.transformer
        ctx = self.make_context(B, S, input_ids.device)
        if position_ids is None:
            position_ids = torch.arange(S, device=input_ids.device).repeat(B)
        flat_ids, flat_pos = input_ids.reshape(-1), position_ids.reshape(-1)
        pdrop, training = self.config.dropout, self.training
        h = self._run(0, lambda: ops.dropout(ops.embedding(flat_ids, t.wte.weight)
                                             + ops.embedding(flat_pos, t.wpe.weight), pdrop, training))   # embd_pdrop
        delta = None
        groups = self._groups()
        for gi, g in enumerate(groups):
            def run_group(a, b=None, g=g):
                for j in g:
                    a, b = t.h[j](a, b, ctx)
                return a, b

            h, delta = self._run(gi + 1, run_group, h) if delta is None else self._run(gi + 1, run_group, h, delta)
        L = len(groups)
        shifted = ops.shift_labels(labels).reshape(-1) if labels is not None else None

        def head(a, b=None):
            x, _ = t.ln_f(b, a) if b is not None else t.ln_f(a)
            if shifted is None:
                return ops.linear(x, self.lm_head.weight)
            return ops.linear_cross_entropy(x, self.lm_head.weight, shifted)

        out = self._run(L + 1, head, h, delta) if delta is not None else self._run(L + 1, head, h)
        if labels is None:
            return CausalLMOutput(logits=out.view(B, S, -1))
        return CausalLMOutput(loss=out)

    def hf_state_dict(self):
        sd = {n: p.data for n, p in self.named_parameters() if n != "lm_head.weight"}
        sd["lm_head.weight"] = self.transformer.wte.weight.data
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
