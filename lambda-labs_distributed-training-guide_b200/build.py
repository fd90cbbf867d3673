"""In-tree build of the sm_100a kernel library (``_lib/libb200train.so``).

The reference has no native code; every kernel it reaches through PyTorch / flash-attn
is replaced here by the sources under ``csrc/``. nvcc cross-compiles without a GPU, so this runs on the
CPU build box; the resulting ``.so`` is git-ignored but travels with the tree.

Usage: ``python -m b200train.build [--force]``
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
LIB_DIR = PKG_DIR / "_lib"
LIB_PATH = LIB_DIR / "libb200train.so"
OBJ_DIR = LIB_DIR / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fopenmp",
    "-I", str(INCLUDE), "-I", str(CSRC),
]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; the b200train kernels need CUDA 12.9+ to build")
    return cand


def _sources() -> list[Path]:
    # *.cu: device + host code through nvcc; *.cpp: host-only code (CPU optimizer) straight through g++
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp"))


CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp"]


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    # the flags without the include directories: those are absolute paths of this checkout, and a tree copied to
    # another machine (gpurun) must still recognise its own up-to-date objects
    h.update(" ".join(f for f in NVCC_FLAGS if not f.startswith("/")).encode())
    return h.hexdigest()


def _src_digest(src: Path, headers) -> str:
    return _digest([src, *headers])


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libb200train.so``. Incremental per file.
    Concurrent callers (the ranks of one job) are serialised by an exclusive lock on ``_lib/.lock``; the
    ones that waited find everything up to date."""
    LIB_DIR.mkdir(exist_ok=True)
    OBJ_DIR.mkdir(exist_ok=True)
    import fcntl

    with open(LIB_DIR / ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            return _build_locked(force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(force: bool, verbose: bool) -> Path:
    nvcc = _nvcc()
    headers = sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h"))
    srcs = _sources()
    todo = []
    for src in srcs:
        obj = OBJ_DIR / (src.stem + ".o")
        stamp = OBJ_DIR / (src.stem + ".sha")
        dig = _src_digest(src, headers)
        if force or not obj.exists() or not stamp.exists() or stamp.read_text() != dig:
            todo.append((src, obj, stamp, dig))

    def compile_one(item):
        src, obj, stamp, dig = item
        if src.suffix == ".cpp":
            cmd = [shutil.which("g++") or "g++", *CXX_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src),
                   "-o", str(obj)]
        else:
            cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"compiler failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        stamp.write_text(dig)
        return src.name

    if todo:
        with ThreadPoolExecutor(max_workers=min(len(todo), os.cpu_count() or 4)) as ex:
            for name in ex.map(compile_one, todo):
                print(f"[b200train.build] compiled {name}", file=sys.stderr)
    link_stamp = OBJ_DIR / "link.sha"
    link_dig = _digest([OBJ_DIR / (s.stem + ".sha") for s in srcs])
    if todo or not LIB_PATH.exists() or not link_stamp.exists() or link_stamp.read_text() != link_dig:
        objs = [str(OBJ_DIR / (s.stem + ".o")) for s in srcs]
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fopenmp",
               "-o", str(LIB_PATH), *objs, "-lgomp"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        link_stamp.write_text(link_dig)
        print(f"[b200train.build] linked {LIB_PATH}", file=sys.stderr)
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB_PATH)
