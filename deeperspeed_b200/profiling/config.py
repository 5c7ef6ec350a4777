"""Flops-profiler sub-config under the reference's import path (deepspeed/profiling/config.py)."""
from ..runtime.config import DeepSpeedFlopsProfilerConfig  # noqa: F401
