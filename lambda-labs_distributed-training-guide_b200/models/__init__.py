"""Model families of the reference guide: GPT-2 (chapters 01/02) and Llama (chapters 04-07, 00)."""
from __future__ import annotations

import torch

from .config import PRESETS, ModelConfig, load_config, to_hf_dict
from .gpt2 import GPT2LMHeadModel
from .llama import CausalLMOutput, LlamaForCausalLM


def build_model(cfg: ModelConfig, dtype=torch.bfloat16, device=None, seed=None, init=True):
    """Random-init causal LM from a config (reference: ``AutoModelForCausalLM.from_config``,
    ``01-single-gpu/train_llm.py:48-49``; meta-device construction ``04-.../train_llm.py:76-79``)."""
    cls = GPT2LMHeadModel if cfg.model_type == "gpt2" else LlamaForCausalLM
    model = cls(cfg, dtype=dtype, device=device)
    if init and (device is None or torch.device(device).type != "meta"):
        model.init_weights(seed)
    model.train()
    return model


__all__ = ["PRESETS", "ModelConfig", "load_config", "to_hf_dict", "GPT2LMHeadModel", "LlamaForCausalLM",
           "CausalLMOutput", "build_model"]
