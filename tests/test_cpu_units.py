"""CPU unit tests of the plumbing: CLI defaults, resume semantics, cosine schedule, packed collator,
timers, metrics keys, cluster monitor parser, config presets, optimizer parity, model parity vs HF."""
import importlib.util
import json
import math
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def test_parser_defaults_match_reference_contract():
    from b200train.cli.train_llm import get_parser

    a = get_parser("01-single-gpu").parse_args(["-e", "x", "-d", "d", "-m", "m"])
    assert (a.save_dir, a.seed, a.num_epochs, a.lr, a.batch_size, a.log_freq, a.ckpt_freq, a.seq_length) == \
        ("../outputs", 0, 100, 3e-5, 1, 100, 500, 1024)
    a = get_parser("04-fsdp").parse_args(["-e", "x", "-d", "d", "-m", "m"])
    assert a.numel_to_wrap == 100_000_000 and a.cpu_offload == "off"
    a = get_parser("05-405b").parse_args(["-e", "x", "-d", "d", "-m", "m"])
    assert a.cpu_offload == "on"
    a = get_parser("07-2d-parallel").parse_args(["-e", "x", "-d", "d", "-m", "m"])
    assert a.tp == 8 and a.seq_length is None
    a = get_parser("00-rime").parse_args(["-e", "x"])
    assert a.num_epochs == 1 and a.log_freq == 50 and a.seq_length == 8192
    with pytest.raises(SystemExit):
        get_parser("01-single-gpu").parse_args(["-d", "d", "-m", "m"])


def test_cosine_schedule_matches_closed_form():
    from b200train.utils.schedule import cosine_lr, make_lr_scheduler, scale_lr

    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=3e-5)
    sched = make_lr_scheduler(opt, 3e-5)
    for step in range(1, 1200):
        opt.step()
        sched.step()
        if step in (1, 10, 500, 1000):
            assert math.isclose(sched.get_last_lr()[0], cosine_lr(step, 3e-5), rel_tol=1e-6)
    assert math.isclose(cosine_lr(1000, 3e-5), 3e-7, rel_tol=1e-9)   # eta_min = lr * 1e-2
    assert scale_lr(1e-3, 8, 32) == 4e-3 and math.isclose(scale_lr(1e-3, 8, 32, "sqrt"), 2e-3)


def test_state_json_roundtrip_and_resume_replay(tmp_path):
    from b200train.cli.train_llm import main

    common = ["-e", "r", "-d", "synthetic:32", "-m", "llama-tiny", "--save-dir", str(tmp_path), "--device", "cpu",
              "-b", "2", "-s", "32", "--log-freq", "2", "--ckpt-freq", "4", "--num-epochs", "1", "--wandb-mode", "off"]
    s1 = main("01-single-gpu", common + ["--max-steps", "6"])
    saved = json.loads((tmp_path / "r" / "state.json").read_text())
    assert set(saved) == {"epoch", "global_step", "epoch_step", "running_loss"}
    assert saved["global_step"] == 4 and saved["epoch_step"] == 4       # last checkpoint, not last step
    for f in ("model.pt", "optimizer.pt", "lr_scheduler.pt", "metrics.jsonl"):
        assert (tmp_path / "r" / f).exists()
    s2 = main("01-single-gpu", common + ["--max-steps", "3"])            # resumes at 4, replays the sampler
    assert s2["global_step"] == 7 and s2["epoch_step"] == 7
    rec = json.loads((tmp_path / "r" / "metrics.jsonl").read_text().splitlines()[0])
    for k in ("global_step", "lr", "running_loss", "epoch", "epoch_progress", "num_batches_remaining", "total_gb",
              "curr_alloc_gb", "peak_alloc_gb", "curr_resv_gb", "peak_resv_gb", "tok/s", "time/total", "time/data",
              "time/forward", "time/backward", "time/update"):
        assert k in rec, k


def test_resume_continues_identically(tmp_path):
    """6 straight steps == 4 steps + checkpoint + resume + 2 steps (model, optimizer, scheduler, data order)."""
    from b200train.cli.train_llm import main

    def args(name, extra):
        return ["-e", name, "-d", "synthetic:32", "-m", "llama-tiny", "--save-dir", str(tmp_path), "--device", "cpu",
                "-b", "2", "-s", "32", "--log-freq", "1", "--ckpt-freq", "4", "--num-epochs", "1", "--wandb-mode", "off",
                "--lr", "1e-3"] + extra

    main("01-single-gpu", args("a", ["--max-steps", "6"]))
    main("01-single-gpu", args("b", ["--max-steps", "4"]))
    main("01-single-gpu", args("b", ["--max-steps", "2"]))
    la = [json.loads(l)["running_loss"] for l in (tmp_path / "a" / "metrics.jsonl").read_text().splitlines()]
    lb = [json.loads(l)["running_loss"] for l in (tmp_path / "b" / "metrics.jsonl").read_text().splitlines()]
    assert len(la) == len(lb) == 6
    for x, y in zip(la, lb):
        assert abs(x - y) < 1e-5, (la, lb)


def test_sharded_resume_restores_learning_rate(tmp_path):
    """DCP (sharded) resume: the optimizer's lr must come back from the checkpoint, otherwise the cosine
    schedule (recursive in ``param_groups[0]['lr']``) continues from the constructor's value."""
    from b200train.cli.train_llm import main

    def args(name, extra):
        return ["-e", name, "-d", "synthetic:32", "-m", "llama-tiny", "--save-dir", str(tmp_path), "--device", "cpu",
                "-b", "2", "-s", "32", "--log-freq", "1", "--ckpt-freq", "4", "--num-epochs", "1", "--wandb-mode", "off",
                "--lr", "1e-3"] + extra

    main("04-fsdp", args("a", ["--max-steps", "6"]))
    main("04-fsdp", args("b", ["--max-steps", "4"]))
    main("04-fsdp", args("b", ["--max-steps", "2"]))
    ra = [json.loads(l) for l in (tmp_path / "a" / "metrics.jsonl").read_text().splitlines()]
    rb = [json.loads(l) for l in (tmp_path / "b" / "metrics.jsonl").read_text().splitlines()]
    assert len(ra) == len(rb) == 6
    for x, y in zip(ra, rb):
        assert math.isclose(x["lr"], y["lr"], rel_tol=1e-9), ([r["lr"] for r in ra], [r["lr"] for r in rb])
        assert abs(x["running_loss"] - y["running_loss"]) < 1e-5


def test_packed_collator_positions_and_mask():
    from b200train.data import block_diagonal_causal_mask, cu_seqlens_from_positions, packed_positions, pad_and_collate

    eos = 128_262
    x = torch.tensor([5, 6, eos, 7, 8, 9, eos, 1, 2])
    keep = x.clone()
    pos, lengths = packed_positions(x)
    assert pos.tolist() == [0, 1, 2, 0, 1, 2, 3, 0, 1] and lengths.tolist() == [3, 4, 2]
    assert torch.equal(x, keep)                                   # the caller's row is not mutated
    mask = block_diagonal_causal_mask(lengths, torch.float32)
    assert mask[4, 3] == 0 and mask[4, 2] == float("-inf") and mask[3, 4] == float("-inf")
    batch = pad_and_collate([{"input_ids": x}, {"input_ids": x}])
    assert batch["position_ids"].shape == (2, 9) and torch.equal(batch["labels"], batch["input_ids"])
    assert cu_seqlens_from_positions(batch["position_ids"]).tolist() == [0, 3, 7, 9, 12, 16, 18]


def test_group_texts_and_seq_length_clamp():
    from b200train.data import group_texts, resolve_seq_length

    out = group_texts({"input_ids": [[1, 2, 3], [4, 5], [6, 7, 8, 9]]}, 4)
    assert out["input_ids"] == [[1, 2, 3, 4], [5, 6, 7, 8]] and out["labels"] == out["input_ids"]
    assert resolve_seq_length(None, 2048, 4096) == 2048
    assert resolve_seq_length(8192, 2048, 4096) == 1024           # > max positions -> min(1024, max)
    assert resolve_seq_length(8192, 2048, 512) == 512


def test_host_adamw_kernel_matches_reference_math():
    """The vectorised host AdamW (csrc/cpu_adamw.cpp: AVX-512 where available, scalar otherwise) against the
    plain PyTorch expression on bf16 buffers, including a length that is not a multiple of the vector width."""
    from b200train.ops.optim import _cpu_kernel_ok, adamw_step, adamw_step_ref

    if not _cpu_kernel_ok():
        pytest.skip("native library unavailable")
    torch.manual_seed(0)
    n = 100_003
    p = (torch.randn(n) * 0.02).bfloat16()
    g = (torch.randn(n) * 0.01).bfloat16()
    m, v = torch.zeros(n).bfloat16(), torch.zeros(n).bfloat16()
    p2, m2, v2 = p.clone(), m.clone(), v.clone()
    for step in (1, 2, 3):
        adamw_step(p, g, m, v, 1e-3, step=step)
        adamw_step_ref(p2, g, m2, v2, 1e-3, 0.9, 0.999, 1e-8, 0.01, step)
    # fused multiply-adds round differently from separate operations: at most one bf16 ulp apart
    assert (p.float() - p2.float()).abs().max().item() <= 2 ** -7 * 0.1
    assert (m.float() - m2.float()).abs().max().item() <= 1e-4
    assert (v.float() - v2.float()).abs().max().item() <= 1e-6


def test_tokenizer_failure_is_loud(tmp_path, monkeypatch):
    """A model name without tokenizer files must raise (as AutoTokenizer does in the reference), not fall
    back to byte-level tokens; the byte tokenizer is for the tiny presets or an explicit opt-in."""
    from b200train.data.datasets import ByteTokenizer, _load_tokenizer, load_and_preprocess_data
    from b200train.models import load_config

    monkeypatch.delenv("B200TRAIN_BYTE_TOKENIZER", raising=False)
    with pytest.raises(RuntimeError, match="cannot load a tokenizer"):
        _load_tokenizer(str(tmp_path / "no-such-model"))
    assert isinstance(_load_tokenizer("llama-tiny"), ByteTokenizer)
    monkeypatch.setenv("B200TRAIN_BYTE_TOKENIZER", "1")
    assert isinstance(_load_tokenizer(str(tmp_path / "no-such-model")), ByteTokenizer)
    text = tmp_path / "corpus.txt"
    text.write_text("hello world " * 200)
    ds = load_and_preprocess_data(str(text), "llama-tiny", load_config("llama-tiny"), seq_length=32)
    assert len(ds) > 0 and ds[0]["input_ids"].shape == (32,)


def test_local_timer_cpu_branch_and_mem_stats():
    from b200train.obs import LocalTimer, get_mem_stats, make_timers

    t = LocalTimer(torch.device("cpu"))
    with t:
        sum(range(1000))
    assert len(t.measurements) == 1 and t.avg_elapsed_ms() >= 0
    t.reset()
    assert t.measurements == []
    assert set(make_timers("cpu", with_waiting=True)) == {"data", "forward", "backward", "update", "waiting"}
    assert set(get_mem_stats("cpu")) == {"total_gb", "curr_alloc_gb", "peak_alloc_gb", "curr_resv_gb", "peak_resv_gb"}


def test_top_cluster_parser():
    spec = importlib.util.spec_from_file_location("top_cluster", ROOT / "top-cluster.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    text = "100, 690.5, 700.0, 78000, 81559\n98, 700.0, 700.0, 79000, 81559\n1234\n1235\n"
    s = mod.parse_node_output(text)
    assert s["ngpus"] == 2 and s["nprocs"] == 2 and abs(s["util"] - 99) < 1e-6 and s["power"] > 98
    table = mod.render({"a": s, "b": {"error": "timeout"}})
    assert "cluster" in table and "unreachable" in table


def test_config_presets_match_hf_parameter_counts():
    from b200train.models import load_config

    assert load_config("gpt2").num_parameters() == 124_439_808
    assert load_config("meta-llama/Llama-3.1-8B").num_parameters() == 8_030_261_248
    assert load_config("llama-3.1-70b").num_parameters() == 70_553_706_496
    assert load_config("llama-3.1-405b").num_parameters() == 405_853_388_800
    assert abs(load_config("llama-3.1-8b").flops_per_token(4096) / 1e9 - 51.47) < 0.05


def test_adamw_reference_matches_torch():
    from b200train.ops.optim import adamw_step

    p = torch.randn(1000)
    g = torch.randn(1000) * 0.01
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in range(1, 4):
        ref.grad = g.clone()
        opt.step()
        adamw_step(p, g, m, v, 1e-3, step=step)
    assert torch.allclose(p, ref.detach(), atol=1e-6)


@pytest.mark.parametrize("name", ["llama-tiny", "gpt2-tiny"])
def test_model_matches_hf_implementation(name):
    """Same weights, same tokens -> same loss as the HF module the reference instantiates."""
    transformers = pytest.importorskip("transformers")
    from b200train.models import build_model, load_config, to_hf_dict

    cfg = load_config(name)
    ours = build_model(cfg, dtype=torch.float32, device="cpu", seed=0)
    hf_cfg = transformers.AutoConfig.for_model(**to_hf_dict(cfg))
    hf_cfg._attn_implementation = "eager"
    hf = transformers.AutoModelForCausalLM.from_config(hf_cfg).float().eval()
    sd = {k: v.clone() for k, v in ours.hf_state_dict().items()}
    missing, unexpected = hf.load_state_dict(sd, strict=False)
    assert not [m for m in missing if "inv_freq" not in m and "attn.bias" not in m and "masked_bias" not in m], missing
    ids = torch.randint(0, cfg.vocab_size, (2, 48))
    with torch.no_grad():
        a = ours(input_ids=ids, labels=ids).loss.item()
        b = hf(input_ids=ids, labels=ids).loss.item()
    assert abs(a - b) < 2e-4, (a, b)


def test_elastic_toy_restarts_and_resumes(tmp_path):
    """torchrun --max-restarts: a rank fails once at step 3, the gang restarts and finishes from step 3.
    (gloo's TCP mesh set-up is occasionally refused right after a restart on this box, so the scenario
    is retried from scratch a few times before it counts as a failure.)"""
    import socket

    last = None
    transport_failures = 0
    attempts = 6
    for attempt in range(attempts):
        sock = socket.socket()
        sock.bind(("127.0.0.1", 0))
        port = sock.getsockname()[1]
        sock.close()
        work = tmp_path / f"try{attempt}"
        work.mkdir()
        state = work / "toy-state.json"
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--max-restarts=4",
               "--monitor-interval", "1", "--master-addr", "127.0.0.1", "--master-port", str(port),
               "--log-dir", str(work / "logs"),
               str(ROOT / "related-topics" / "elastic-training" / "toy.py"), "--state-file", str(state),
               "--total-steps", "6", "--fail-at-step", "3", "--fail-rank", "1"]
        env = {**os.environ, "OMP_NUM_THREADS": "1", "GLOO_SOCKET_IFNAME": "lo"}
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=150, env=env)
        except subprocess.TimeoutExpired as e:
            last = f"timeout: {e}"
            transport_failures += 1
            continue
        out = r.stdout + r.stderr
        if r.returncode == 0 and state.exists() and json.loads(state.read_text())["num_steps"] == 6 \
                and "resuming at step 3" in out and "injected failure" in out:
            return
        last = out[-1500:]
        if "connectFullMesh" in out or "Connection refused" in out or "timed out connecting" in out:
            transport_failures += 1
    if transport_failures == attempts:
        # every attempt died inside gloo's TCP mesh set-up (seen on loaded CI containers): the restart / resume
        # logic under test never got to run, which is not a verdict on it
        pytest.skip(f"gloo could not build its TCP mesh in {attempts} attempts on this machine: {last[-300:]}")
    pytest.fail(f"elastic toy never completed: {last}")


def test_pretrained_roundtrip_and_vocab_growth(tmp_path):
    """save_pretrained -> load_pretrained into a differently-initialised FSDP engine reproduces the loss;
    loading into a model with added tokens fills the new rows with the mean embedding."""
    from b200train.models import load_config
    from b200train.models.pretrained import load_pretrained, save_pretrained
    from b200train.parallel import make_strategy

    cfg = load_config("llama-tiny")
    a = make_strategy("single", cfg, "cpu", dtype=torch.float32, seed=1)
    save_pretrained(a, tmp_path / "ckpt")
    ids = torch.randint(0, cfg.vocab_size, (2, 32))
    ref = a.forward({"input_ids": ids, "labels": ids}).loss.item()
    b = make_strategy("fsdp", cfg, "cpu", dtype=torch.float32, seed=2)
    assert load_pretrained(b, tmp_path / "ckpt") > 0
    assert abs(b.forward({"input_ids": ids, "labels": ids}).loss.item() - ref) < 1e-5
    big = load_config("llama-tiny")
    big.vocab_size += 16
    c = make_strategy("single", big, "cpu", dtype=torch.float32, seed=3)
    load_pretrained(c, tmp_path / "ckpt")
    emb = c.model.model.embed_tokens.weight
    assert torch.allclose(emb[-1], a.model.model.embed_tokens.weight.mean(0), atol=1e-6)
    assert torch.equal(emb[:cfg.vocab_size], a.model.model.embed_tokens.weight)


def test_chunked_linear_cross_entropy_matches_full():
    """LM head + CE over row chunks (``--ce-chunk-rows``): loss and both gradients equal the all-rows reference
    for chunk sizes that divide the rows, do not divide them, and exceed them; ignored rows contribute nothing."""
    from b200train.ops.cross_entropy import IGNORE, linear_cross_entropy, linear_cross_entropy_ref

    torch.manual_seed(0)
    T, H, V = 37, 16, 101
    x = torch.randn(T, H, requires_grad=True)
    w = (torch.randn(V, H) * 0.3).requires_grad_()
    labels = torch.randint(0, V, (T,))
    labels[::5] = IGNORE
    ref = linear_cross_entropy_ref(x, w, labels)
    gx, gw = torch.autograd.grad(ref * 1.7, (x, w))
    for chunk in (1, 8, 10, 36, 37, 64):
        loss = linear_cross_entropy(x, w, labels, chunk_rows=chunk)
        cx, cw = torch.autograd.grad(loss * 1.7, (x, w))
        assert abs(loss.item() - ref.item()) < 1e-6
        assert (cx - gx).abs().max().item() < 1e-6 and (cw - gw).abs().max().item() < 1e-6
        assert cx[::5].abs().max().item() == 0.0


def test_warmup_cosine_schedule_of_the_deepspeed_variant():
    """``WarmupCosineLR`` as configured by the reference's ``ds_config.json``: linear warm-up, half a cosine down
    to ``cos_min_ratio`` at ``total_num_steps``, flat afterwards; as a torch scheduler it drives ``param_groups``."""
    from b200train.utils.schedule import make_warmup_cosine, warmup_cosine_ratio

    kw = dict(total_num_steps=100, warmup_num_steps=10, cos_min_ratio=1e-2)
    assert warmup_cosine_ratio(0, **kw) == 0.0
    assert math.isclose(warmup_cosine_ratio(5, **kw), 0.5)
    assert math.isclose(warmup_cosine_ratio(10, **kw), 1.0)
    assert math.isclose(warmup_cosine_ratio(55, **kw), 1e-2 + 0.99 * 0.5)
    assert math.isclose(warmup_cosine_ratio(100, **kw), 1e-2)
    assert math.isclose(warmup_cosine_ratio(500, **kw), 1e-2)
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=3e-5)
    sched = make_warmup_cosine(opt, total_num_steps=1000, warmup_num_steps=0, cos_min_ratio=1e-2, extra="ignored")
    assert math.isclose(opt.param_groups[0]["lr"], 3e-5)
    for _ in range(1000):
        opt.step()
        sched.step()
    assert math.isclose(opt.param_groups[0]["lr"], 3e-7, rel_tol=1e-6)


def _philox_np(ctr, k0, k1):
    import numpy as np

    c = [np.uint32(v) for v in ctr]
    for _ in range(10):
        p0 = np.uint64(0xD2511F53) * np.uint64(c[0])
        p1 = np.uint64(0xCD9E8D57) * np.uint64(c[2])
        hi0, lo0 = np.uint32(p0 >> np.uint64(32)), np.uint32(p0 & np.uint64(0xFFFFFFFF))
        hi1, lo1 = np.uint32(p1 >> np.uint64(32)), np.uint32(p1 & np.uint64(0xFFFFFFFF))
        c = [hi1 ^ c[1] ^ np.uint32(k0), lo1, hi0 ^ c[3] ^ np.uint32(k1), lo0]
        k0, k1 = (k0 + 0x9E3779B9) & 0xFFFFFFFF, (k1 + 0xBB67AE85) & 0xFFFFFFFF
    return [int(v) for v in c]


def test_host_dropout_is_philox_and_regenerates_its_mask_in_backward():
    """``csrc/cpu_dropout.cpp`` against an independent Philox4x32-10 in numpy with the device kernel's counter
    layout (vector of 8 elements -> two blocks, keep iff draw >= p * 2^32): the AVX-512 groups and the scalar
    tail give the same mask; backward applies the same mask to the gradient without having stored it."""
    from b200train.ops.dropout import _cpu_kernel_ok, dropout

    if not _cpu_kernel_ok():
        pytest.skip("native library not built")
    n, p = 128 * 3 + 29, 0.25
    seed, offset = 0x0123456789ABCDEF, 0x0FEDCBA987654321
    k0 = (seed ^ offset) & 0xFFFFFFFF
    k1 = ((seed >> 32) ^ (offset >> 32) ^ 0x27D4EB2F) & 0xFFFFFFFF
    keep = torch.zeros(n, dtype=torch.bool)
    for i in range((n + 7) // 8):
        r = _philox_np([i, 0, 0, 0], k0, k1) + _philox_np([i, 0, 1, 0], k0, k1)
        for j in range(8):
            if i * 8 + j < n:
                keep[i * 8 + j] = r[j] >= int(p * 2 ** 32)
    assert 0.6 < keep.float().mean().item() < 0.9
    sc = torch.tensor(1.0 / (1.0 - p), dtype=torch.float32)   # the kernel multiplies by the fp32 reciprocal
    for dt in (torch.float32, torch.bfloat16):
        x = (torch.arange(1, n + 1, dtype=torch.float32) / 64).to(dt).requires_grad_()
        y = dropout(x, p, philox=(seed, offset))
        want = torch.where(keep, (x.detach().float() * sc).to(dt), torch.zeros(n, dtype=dt))
        assert torch.equal(y.detach(), want), dt
        g = torch.ones(n, dtype=dt)
        y.backward(g)
        assert torch.equal(x.grad, torch.where(keep, (g.float() * sc).to(dt), torch.zeros(n, dtype=dt))), dt
    # draws from the default generator: repeatable under manual_seed, different between calls
    torch.manual_seed(5)
    a, b = dropout(torch.ones(4096), 0.5), dropout(torch.ones(4096), 0.5)
    torch.manual_seed(5)
    a2 = dropout(torch.ones(4096), 0.5)
    assert torch.equal(a, a2) and not torch.equal(a, b) and abs(a.mean().item() - 1.0) < 0.1


def test_local_corpus_is_tokenized_in_batches_and_cached(tmp_path, monkeypatch):
    """A local .txt / .jsonl corpus: concatenated token stream cut into blocks, cached on disk (second call does
    not tokenize), cache invalidated when a file changes."""
    from b200train.data import datasets as D
    from b200train.models import load_config

    monkeypatch.setenv("B200TRAIN_CACHE", str(tmp_path / "cache"))
    corpus = tmp_path / "corpus"
    corpus.mkdir()
    (corpus / "a.txt").write_text("hello world " * 40)
    (corpus / "b.jsonl").write_text("\n".join(json.dumps({"text": f"row {i} " * 10}) for i in range(30)))
    cfg = load_config("llama-tiny")
    calls = []
    real = D.ByteTokenizer.__call__
    monkeypatch.setattr(D.ByteTokenizer, "__call__", lambda self, texts: (calls.append(len(texts)), real(self, texts))[1])
    ds = D.load_and_preprocess_data(str(corpus), "llama-tiny", cfg, 32)
    assert sum(calls) == 31 and len(calls) == 1                      # one batch for all 31 texts
    want = ("hello world " * 40).encode() + b""
    assert bytes(ds[0]["input_ids"].tolist()) == want[:32] and torch.equal(ds[0]["labels"], ds[0]["input_ids"])
    n_blocks = len(ds)
    assert n_blocks == (len(want) + 1 + sum(len((f"row {i} " * 10).encode()) + 1 for i in range(30))) // 32
    calls.clear()
    ds2 = D.load_and_preprocess_data(str(corpus), "llama-tiny", cfg, 32)
    assert calls == [] and len(ds2) == n_blocks and torch.equal(ds2[3]["input_ids"], ds[3]["input_ids"])
    (corpus / "a.txt").write_text("changed " * 50)
    D.load_and_preprocess_data(str(corpus), "llama-tiny", cfg, 32)
    assert calls == [31]                                              # the cache key covers size and mtime


def test_synthetic_packed_rows_hold_several_documents():
    """Chapter 00 / ``--packed`` with synthetic data: rows are documents of 701 tokens separated by the EOS id (which
    appears nowhere else), and the collator restarts the position ids at every document."""
    from functools import partial

    from b200train.data.datasets import SyntheticTokens
    from b200train.data.packed import cu_seqlens_from_positions, pad_and_collate

    eos = 999
    ds = SyntheticTokens(1000, 2048, n_samples=4, seed=1, eos_id=eos)
    row = ds[2]["input_ids"]
    ends = torch.nonzero(row == eos).flatten().tolist()
    assert ends == [700, 1401] and int(row.max()) <= 999
    batch = partial(pad_and_collate, eos_id=eos)([ds[0], ds[1]])
    pos = batch["position_ids"]
    assert pos.shape == (2, 2048) and pos[0, 700] == 700 and pos[0, 701] == 0 and pos[0, 1402] == 0
    assert int(pos[0, -1]) == 2048 - 1402 - 1
    cu = cu_seqlens_from_positions(pos)
    assert cu.tolist() == [0, 701, 1402, 2048, 2749, 3450, 4096]
    plain = SyntheticTokens(1000, 2048, n_samples=4, seed=1)
    assert not torch.equal(plain[2]["input_ids"], row)


def test_cpu_attention_with_dropout_uses_the_philox_mask():
    """The CPU attention path with ``dropout_p`` > 0 (GPT-2 on CPU): explicit products with the host Philox dropout
    on the probabilities. Same generator state -> same result as a hand-written softmax(QK^T)V with ``ops.dropout``
    at the same site; without dropout it equals ATen's fused attention; gradients reach q, k and v."""
    from b200train.ops.attention import attention_ref
    from b200train.ops.dropout import dropout

    torch.manual_seed(0)
    L, H, D = 48, 3, 16
    qkv = torch.randn(2 * L, 3 * H * D, requires_grad=True)
    cu = torch.tensor([0, L, 2 * L], dtype=torch.int32)
    torch.manual_seed(11)
    out = attention_ref(qkv, cu, L, H, H, D, causal=True, dropout_p=0.25)
    q, k, v = (t.reshape(2 * L, H, D) for t in qkv.detach().split(H * D, dim=1))
    torch.manual_seed(11)
    rows = []
    for s in (0, L):
        qi, ki, vi = (t[s:s + L].transpose(0, 1).unsqueeze(0) for t in (q, k, v))
        att = torch.matmul(qi * D ** -0.5, ki.transpose(-1, -2))
        att = att.masked_fill(torch.ones(L, L, dtype=torch.bool).triu_(1), float("-inf"))
        pi = dropout(torch.softmax(att, dim=-1), 0.25)
        rows.append(torch.matmul(pi, vi)[0].transpose(0, 1).reshape(L, H * D))
    assert torch.allclose(out.detach(), torch.cat(rows), atol=1e-6)
    assert not torch.allclose(out.detach(), attention_ref(qkv, cu, L, H, H, D, causal=True).detach(), atol=1e-3)
    out.square().sum().backward()
    g = qkv.grad.reshape(2 * L, 3, H * D)
    assert all(g[:, j].abs().sum().item() > 0 for j in range(3))


def test_reduce_scatter_transport_policy():
    """``PeerComm._transport``: which path a reduce-scatter takes (copy engines + local reduce, in-switch
    ``multimem.ld_reduce``, SM loads from the peers) as a function of world size, fusion with AdamW, the
    availability of a multicast mapping and the ``B200TRAIN_RS_TRANSPORT`` override."""
    from types import SimpleNamespace

    from b200train.parallel.peer import PeerComm

    def pick(world, fused=False, mc=True, override="auto", nvls_flag=True, forced=None):
        stub = SimpleNamespace(rs_transport=override, world=world, nvls=nvls_flag)
        return PeerComm._transport(stub, forced, mc, fused)

    assert pick(2) == "ce" and pick(4) == "nvls" and pick(8) == "nvls"          # plain: in-switch from 4 ranks
    assert pick(4, mc=False) == "ce" and pick(8, nvls_flag=False) == "ce"       # no multicast mapping: copy engines
    assert all(pick(w, fused=True) == "ce" for w in (2, 4, 8))                  # fused with AdamW: always the CE pull
    assert pick(8, override="peer") == "peer" and pick(2, override="nvls") == "nvls"
    assert pick(2, override="nvls", mc=False) == "ce"                           # an override cannot invent a mapping
    assert pick(8, forced=True) == "nvls" and pick(8, forced=True, mc=False) == "peer" and pick(8, forced=False) == "peer"
