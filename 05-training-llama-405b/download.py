#!/usr/bin/env python
"""Pre-download a model's config, tokenizer and (unless ``--skip-model``) weights into ``HF_HOME`` before a
multi-node run, so 64 ranks do not hit the hub at once (reference: ``05-training-llama-405b/download.py``;
node-local disks load 405B in ~3 min instead of ~50 min, ``05-.../README.md:56``).

Needs network access and the ``transformers`` package; the weights are instantiated on the meta device
so only the download happens. On an offline box the script explains what to copy instead.
"""
import argparse
import sys


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model-name", "-m", default="meta-llama/Meta-Llama-3.1-405B")
    ap.add_argument("--skip-model", action="store_true", help="fetch config + tokenizer only")
    args = ap.parse_args(argv)
    try:
        import torch
        from transformers import AutoConfig, AutoModelForCausalLM, AutoTokenizer

        AutoConfig.from_pretrained(args.model_name)
        AutoTokenizer.from_pretrained(args.model_name)
        if not args.skip_model:
            with torch.device("meta"):
                AutoModelForCausalLM.from_pretrained(args.model_name)
        print(f"{args.model_name}: cached under HF_HOME")
    except Exception as e:
        print(f"download failed ({type(e).__name__}: {e}).\nOffline: copy a directory containing config.json "
              f"(and tokenizer files / *.safetensors) to each node and pass it as --model-name.", file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
