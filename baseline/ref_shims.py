"""Environment shims that let the UNMODIFIED reference (installed in baseline/_ref) import on torch 2.x / py3.12.
Nothing under baseline/_ref is edited: missing legacy modules / attributes are provided from the outside.

  torch._six                       removed in torch 2.0   -> stub module (inf, string_classes, container_abcs)
  tensorboardX                     not installed          -> stub SummaryWriter (tensorboard is disabled in the bench)
  collections.Mapping / Sequence   moved to collections.abc in py3.10
  torch.cuda.memory_cached(...)    renamed memory_reserved
  op_builder arch string           `cc[0] + cc[2]` mis-parses "10.0" (reference op_builder/builder.py:306) -> the bench
                                   uses `torch_adam: true` so no CUDA op has to be JIT-built; only the C++ flatten op is.
"""
import collections
import collections.abc
import math
import os
import sys
import types


def install(ref_root):
    if ref_root not in sys.path:
        sys.path.insert(0, ref_root)
    import torch
    if "torch._six" not in sys.modules:
        six = types.ModuleType("torch._six")
        six.inf = math.inf
        six.nan = math.nan
        six.string_classes = (str, bytes)
        six.int_classes = (int,)
        six.container_abcs = collections.abc
        sys.modules["torch._six"] = six
        torch._six = six
    try:
        import tensorboardX  # noqa: F401
    except Exception:
        tbx = types.ModuleType("tensorboardX")

        class SummaryWriter:
            def __init__(self, *a, **k):
                pass

            def add_scalar(self, *a, **k):
                pass

            def flush(self):
                pass

        tbx.SummaryWriter = SummaryWriter
        sys.modules["tensorboardX"] = tbx
    for name in ("Mapping", "Sequence", "MutableMapping", "Iterable", "Callable"):
        if not hasattr(collections, name):
            setattr(collections, name, getattr(collections.abc, name))
    if not hasattr(torch.cuda, "memory_cached"):
        torch.cuda.memory_cached = torch.cuda.memory_reserved
    if not hasattr(torch.cuda, "max_memory_cached"):
        torch.cuda.max_memory_cached = torch.cuda.max_memory_reserved
    import fractions
    if not hasattr(fractions, "gcd"):
        fractions.gcd = math.gcd
    os.environ.setdefault("TORCH_EXTENSIONS_DIR", "/tmp/ref_torch_extensions")
