from .profiler import FlopsProfiler, get_model_profile
