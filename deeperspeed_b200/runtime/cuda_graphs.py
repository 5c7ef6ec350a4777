"""CUDA-graph capture of launch-bound module stacks.

Small models (BERT-large layers at 8 K tokens: ~40 kernels of 10-60 us per layer) are bound by the Python / ctypes
launch path, not by the GPU. `graph_module` captures forward AND backward of a module (or a tuple of modules) into CUDA
graphs through `torch.cuda.make_graphed_callables`, so one replay issues the whole kernel sequence. Every kernel of this
framework is capturable: scratch is caller-owned or comes from the torch allocator (graph-private pool), tensor maps
are encoded on the host at capture time into kernel parameters, and the CTA-pair GEMM's split-K workspace is keyed by
stream and is created during the warm-up iterations that precede the capture on the same side stream.

Restrictions (checked where possible): static shapes; no host synchronisation inside the module; dropout must be off -
the Philox seed/offset are kernel arguments and would be frozen into the graph.
"""
import torch


def _dropout_active(module):
    for m in module.modules():
        p = getattr(m, "p", None)
        if isinstance(m, torch.nn.Dropout) and p and m.training:
            return True
        cfg = getattr(m, "config", None)
        for name in ("hidden_dropout_ratio", "attn_dropout_ratio"):
            if cfg is not None and getattr(cfg, name, 0) and m.training:
                return True
    return False


def graph_module(module, sample_args, num_warmup_iters=3, allow_unused_input=True):
    """Return a drop-in replacement for `module` (or a tuple of modules) whose forward and backward replay CUDA
    graphs. `sample_args`: tuple of example inputs (tuple of tuples for several modules) with the shapes, dtypes and
    requires_grad flags of the real inputs."""
    if not torch.cuda.is_available():
        return module
    mods = module if isinstance(module, (tuple, list)) else (module,)
    for m in mods:
        if _dropout_active(m):
            raise ValueError("graph_module: dropout is active; its Philox offsets would be frozen into the graph")
    return torch.cuda.make_graphed_callables(module if isinstance(module, (tuple, list)) else module, sample_args,
                                             num_warmup_iters=num_warmup_iters,
                                             allow_unused_input=allow_unused_input)
