from . import flops_profiler  # noqa: F401  (deepspeed.profiling.flops_profiler is reachable after a bare import)
