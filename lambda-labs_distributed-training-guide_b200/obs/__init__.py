from . import nvtx
from .metrics import LOGGER, MetricsSink, build_record, setup_logging
from .timers import PHASES, CudaEventTimer, LocalTimer, get_mem_stats, make_timers

__all__ = ["LOGGER", "MetricsSink", "build_record", "setup_logging", "PHASES", "CudaEventTimer", "LocalTimer",
           "get_mem_stats", "make_timers", "nvtx"]
