"""Phase timers and memory statistics.

``LocalTimer`` reproduces the reference's timer (device synchronise, host clock, synchronise again;
CPU branch included): ``01-single-gpu/train_llm.py:262-288``; the four phases ``data / forward /
backward / update`` (``:115``) plus the optional ``waiting`` timer around barriers
(``related-topics/optimizing-data-loading/README.md:57-88``). ``CudaEventTimer`` is the device-side
variant used where phases must be allowed to overlap (it does not synchronise on entry/exit).
"""
from __future__ import annotations

import time

import torch


class LocalTimer:
    def __init__(self, device: torch.device):
        device = torch.device(device)
        if device.type == "cuda":
            self.synchronize = lambda: torch.cuda.synchronize(device=device)
        else:
            self.synchronize = lambda: None
        self.measurements = []
        self.start_time = None

    def __enter__(self):
        self.synchronize()
        self.start_time = time.time()
        return self

    def __exit__(self, exc_type, value, traceback):
        if traceback is None:
            self.synchronize()
            self.measurements.append(time.time() - self.start_time)
        self.start_time = None

    def avg_elapsed_ms(self) -> float:
        if not self.measurements:
            return 0.0
        return 1000 * (sum(self.measurements) / len(self.measurements))

    def reset(self):
        self.measurements = []
        self.start_time = None


class CudaEventTimer:
    """Records CUDA events on the current stream; elapsed times are resolved lazily in
    ``avg_elapsed_ms`` (one synchronise per logging window instead of two per phase)."""

    def __init__(self, device: torch.device):
        self.device = torch.device(device)
        self.pairs = []
        self._cpu = LocalTimer(device) if self.device.type != "cuda" else None

    def __enter__(self):
        if self._cpu is not None:
            return self._cpu.__enter__()
        self._start = torch.cuda.Event(enable_timing=True)
        self._start.record()
        return self

    def __exit__(self, exc_type, value, traceback):
        if self._cpu is not None:
            return self._cpu.__exit__(exc_type, value, traceback)
        if traceback is None:
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            self.pairs.append((self._start, end))

    def avg_elapsed_ms(self) -> float:
        if self._cpu is not None:
            return self._cpu.avg_elapsed_ms()
        if not self.pairs:
            return 0.0
        torch.cuda.synchronize(self.device)
        return sum(s.elapsed_time(e) for s, e in self.pairs) / len(self.pairs)

    def reset(self):
        self.pairs = []
        if self._cpu is not None:
            self._cpu.reset()


PHASES = ["data", "forward", "backward", "update"]


def make_timers(device, kind: str = "local", with_waiting: bool = False):
    cls = LocalTimer if kind == "local" else CudaEventTimer
    names = PHASES + (["waiting"] if with_waiting else [])
    return {k: cls(device) for k in names}


def get_mem_stats(device=None) -> dict:
    """Keys of the reference record (``01-single-gpu/train_llm.py:250-259``); zeros on CPU."""
    device = torch.device(device) if device is not None else None
    if device is None or device.type != "cuda":
        return {"total_gb": 0.0, "curr_alloc_gb": 0.0, "peak_alloc_gb": 0.0, "curr_resv_gb": 0.0,
                "peak_resv_gb": 0.0}
    mem = torch.cuda.memory_stats(device)
    props = torch.cuda.get_device_properties(device)
    return {
        "total_gb": 1e-9 * props.total_memory,
        "curr_alloc_gb": 1e-9 * mem.get("allocated_bytes.all.current", 0),
        "peak_alloc_gb": 1e-9 * mem.get("allocated_bytes.all.peak", 0),
        "curr_resv_gb": 1e-9 * mem.get("reserved_bytes.all.current", 0),
        "peak_resv_gb": 1e-9 * mem.get("reserved_bytes.all.peak", 0),
    }
