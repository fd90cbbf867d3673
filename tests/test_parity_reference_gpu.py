"""Parity with the reference itself on the GPU ("parity harness").

The UNMODIFIED ``01-single-gpu/train_llm.py`` of the reference (HF ``LlamaForCausalLM``, ``AdamW(fused=True)``,
cosine schedule: the copy under ``baseline/_ref``) and this framework's single-device engine are started
from the same weights (written once as HF-named safetensors) and fed the same synthetic batches in the same
order; the loss of each of the first optimizer steps must agree within bf16 noise. A wrong gradient, a
different update rule or a different loss normalisation shows up as a drifting curve."""
import json
import sys
import tempfile
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
STEPS, B, S, LR = 6, 2, 256, 1e-3


def test_first_steps_match_reference_script():
    sys.path.insert(0, str(ROOT / "baseline"))
    import run_reference as rr

    ok, why = rr.reference_available()
    if not ok:
        pytest.skip(why)
    for mod in ("wandb", "datasets", "transformers", "tqdm"):
        pytest.importorskip(mod)
    from b200train.models import load_config, to_hf_dict
    from b200train.models.pretrained import save_pretrained
    from b200train.parallel import make_strategy

    cfg = load_config("llama-smoke")
    eng = make_strategy("single", cfg, "cuda:0", lr=LR, seed=0)
    tmp = Path(tempfile.mkdtemp(prefix="b200train-parity-"))
    save_pretrained(eng, tmp / "init")

    # the reference's loader (in-order in parity mode): item i is seeded with 1234 + i, batch k = items [k*B, (k+1)*B)
    def batch(k):
        rows = []
        for i in range(k * B, (k + 1) * B):
            g = torch.Generator().manual_seed(1234 + i)
            rows.append(torch.randint(0, cfg.vocab_size, (S,), generator=g))
        ids = torch.stack(rows)
        return {"input_ids": ids, "labels": ids}

    ours = [eng.train_step(batch(k)).item() for k in range(STEPS)]
    torch.cuda.synchronize()
    del eng
    res = rr.run_reference("single", to_hf_dict(cfg), B, S, STEPS, 0, extra_args=("--lr", str(LR)),
                           init_weights_dir=tmp / "init", capture_losses=True)
    ref = res["losses"][:STEPS]
    assert len(ref) == STEPS, res["losses"]
    (tmp / "losses.json").write_text(json.dumps({"ours": ours, "reference": ref}))
    print("ours     ", [round(x, 4) for x in ours])
    print("reference", [round(x, 4) for x in ref])
    assert abs(ours[0] - ref[0]) < 0.02, (ours, ref)          # same weights, same batch: the forward agrees
    for a, b in zip(ours, ref):
        assert abs(a - b) < 0.05, (ours, ref)
    assert ref[-1] < ref[0] and ours[-1] < ours[0]
