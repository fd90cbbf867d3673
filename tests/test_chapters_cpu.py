"""Chapter entry points end to end on CPU (gloo): each script must train a few steps with the reference's
flags, write the reference's on-disk layout, and resume from it."""
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(script, nproc, args, timeout=420):
    env = {**os.environ, "OMP_NUM_THREADS": "1", "WANDB_MODE": "disabled", "GLOO_SOCKET_IFNAME": "lo"}
    if nproc == 1:
        cmd = [sys.executable, str(ROOT / script)] + args
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
               "--master-addr", "127.0.0.1", "--master-port", str(_port()), str(ROOT / script)] + args
    for attempt in range(3):
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=str(ROOT))
        out = r.stdout + r.stderr
        # gloo's TCP mesh set-up occasionally fails on a loaded box before the script has done anything: retry
        # that (on a fresh port), and only that
        if r.returncode != 0 and attempt < 2 and nproc > 1 and "global_step" not in out and any(
                t in out for t in ("connectFullMesh", "Connection refused", "Connection reset by peer")):
            cmd = [c if not c.isdigit() or cmd[i - 1] != "--master-port" else str(_port()) for i, c in enumerate(cmd)]
            continue
        break
    assert r.returncode == 0, out[-3000:]
    return out


COMMON = ["-d", "synthetic:64", "-m", "llama-tiny", "--device", "cpu", "-b", "2", "-s", "32", "--log-freq", "2",
          "--ckpt-freq", "2", "--num-epochs", "1", "--wandb-mode", "off"]


@pytest.mark.parametrize("script,nproc,extra,sharded", [
    ("01-single-gpu/train_llm.py", 1, [], False),
    ("02-distributed-data-parallel/train_llm.py", 2, [], True),
    ("04-fully-sharded-data-parallel/train_llm.py", 2, ["--cpu-offload", "off", "--numel-to-wrap", "1000"], True),
    ("04-fully-sharded-data-parallel/train_llm.py", 4, ["--sharding-strategy", "hybrid_shard", "--replicas", "2"], True),
    ("04-fully-sharded-data-parallel/train_llm.py", 2, ["--sharding-strategy", "shard_grad_op"], True),
    ("05-training-llama-405b/train_llm.py", 2, ["--num-workers", "0"], True),
    ("06-tensor-parallel/train_llm.py", 2, ["--num-workers", "0"], True),
    ("07-2d-parallel/train_llm.py", 4, ["--tp", "2", "--num-workers", "0"], True),
])
def test_chapter_trains_checkpoints_and_resumes(tmp_path, script, nproc, extra, sharded):
    args = ["-e", "exp", "--save-dir", str(tmp_path)] + COMMON + extra
    _run(script, nproc, args + ["--max-steps", "3"])
    exp = tmp_path / "exp"
    state = json.loads((exp / "state.json").read_text())
    assert state["global_step"] == 2 and set(state) == {"epoch", "global_step", "epoch_step", "running_loss"}
    assert (exp / "lr_scheduler.pt").exists()
    if sharded:
        files = {p.name for p in (exp / "checkpoint").iterdir()}
        assert ".metadata" in files and any(f.endswith(".distcp") for f in files), files
        assert (exp / "rank-0").is_dir()
    else:
        assert (exp / "model.pt").exists() and (exp / "optimizer.pt").exists()
    out = _run(script, nproc, args + ["--max-steps", "2"])
    assert "Resumed=True" in out
    assert json.loads((exp / "state.json").read_text())["global_step"] == 4


def test_packed_chapter_and_deepspeed_entry(tmp_path):
    out = _run("00-rime/train_llm.py", 1, ["-e", "rime", "--save-dir", str(tmp_path), "-m", "llama-tiny", "-d",
                                           "synthetic:16", "--device", "cpu", "-b", "1", "-s", "64", "--log-freq", "2",
                                           "--max-steps", "2", "--wandb-mode", "off", "--add-tokens", "0"])
    assert "global_step" in out
    out = _run("alternative-frameworks/deepspeed/train_llm.py", 2,
               ["-e", "ds", "--save-dir", str(tmp_path), "-m", "llama-tiny", "-d", "synthetic:32", "--device", "cpu",
                "-s", "32", "--log-freq", "1", "--max-steps", "2", "--wandb-mode", "off", "--deepspeed",
                "--deepspeed_config", str(ROOT / "alternative-frameworks/deepspeed/ds_config.json")])
    assert "global_step" in out
