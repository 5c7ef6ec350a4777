"""FlopsProfiler: MACs, parameters and forward latency of every module of a model.

Interface and report layout per reference deepspeed/profiling/flops_profiler/profiler.py:11-868 (`FlopsProfiler`
start / stop / reset / end, `get_total_*`, `print_model_profile`, `print_model_aggregated_profile`,
`get_model_profile`, the `*_to_string` helpers). Own mechanism:
  * the profiler keeps one record per module in a dictionary (nothing is stored on the modules), with a stack of
    "currently executing" records that the patched operators charge their MACs to;
  * operator rules are a table {namespace: {function: MAC formula}} that also covers this framework's fused kernels
    (`ops.kernels.linear / linear_gelu / linear_residual / mlp_residual`, the flash-attention block), which never pass
    through torch.nn.functional and would otherwise count as zero;
  * on CUDA the latency of a module is measured with a pair of events and read once, when the numbers are asked for,
    instead of synchronising the device before and after every module.
"""
import time
from functools import wraps

import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------- MAC formulas
def _prod(dims):
    out = 1
    for v in dims:
        out *= int(v)
    return out


def _linear_macs(input, weight, bias=None, *args, **kwargs):
    return input.numel() * weight.shape[0]


def _per_element(input, *args, **kwargs):
    return input.numel()


def _norm_macs(input, *args, **kwargs):
    return 5 * input.numel()


def _zero(*args, **kwargs):
    return 0


def _conv_macs(input, weight, bias=None, stride=1, padding=0, dilation=1, groups=1):
    spatial = list(input.shape[2:])
    kernel = list(weight.shape[2:])
    as_tuple = lambda v: v if isinstance(v, tuple) else (v,) * len(spatial)
    out_positions = input.shape[0]
    for size, k, s, p, d in zip(spatial, kernel, as_tuple(stride), as_tuple(padding), as_tuple(dilation)):
        p = 0 if isinstance(p, str) else p
        out_positions *= (size + 2 * p - (d * (k - 1) + 1)) // s + 1
    per_position = _prod(kernel) * input.shape[1] * (weight.shape[0] // groups)
    return int(out_positions * (per_position + (weight.shape[0] if bias is not None else 0)))


def _matmul_macs(input, other, *args, **kwargs):
    return input.numel() * other.shape[-1]


def _sdpa_macs(q, k, v, *args, **kwargs):
    return 2 * q.numel() * k.shape[-2]          # scores + context


def _linear_residual_macs(x, w, b, r):
    return x.numel() * w.shape[0]


def _mlp_residual_macs(x, w1, b1, w2, b2, r):
    rows = x.numel() // x.shape[-1]
    return rows * (w1.shape[0] * w1.shape[1] + w2.shape[0] * w2.shape[1])


def _causal_attention_macs(qkv, H, D, rot_dim, base=10000.0):
    B, S = qkv.shape[0], qkv.shape[1]
    return 2 * B * H * S * S * D // 2           # causal: half of the score / context products


_TORCH_RULES = {
    "linear": _linear_macs, "conv1d": _conv_macs, "conv2d": _conv_macs, "conv3d": _conv_macs,
    "relu": _per_element, "gelu": _per_element, "silu": _per_element, "elu": _per_element, "leaky_relu": _per_element,
    "relu6": _per_element, "prelu": _per_element, "softmax": _per_element,
    "batch_norm": _norm_macs, "layer_norm": _norm_macs, "instance_norm": _norm_macs, "group_norm": _norm_macs,
    "avg_pool1d": _per_element, "avg_pool2d": _per_element, "avg_pool3d": _per_element, "max_pool1d": _per_element,
    "max_pool2d": _per_element, "max_pool3d": _per_element, "adaptive_avg_pool2d": _per_element,
    "adaptive_max_pool2d": _per_element, "embedding": _zero, "scaled_dot_product_attention": _sdpa_macs,
}
_KERNEL_RULES = {"linear": _linear_macs, "linear_gelu": _linear_macs, "linear_residual": _linear_residual_macs,
                 "mlp_residual": _mlp_residual_macs}
_ATTENTION_RULES = {"causal_self_attention": _causal_attention_macs}

# kept for code that extends the profiler the reference's way (module type -> forward hook)
MODULE_HOOK_MAPPING = {}
module_flop_count = []          # the live stack of MAC lists (one per executing module)
old_functions = {}
_depth = [0]


def _namespaces():
    spaces = [(F, _TORCH_RULES)]
    try:
        from ...ops import kernels as K
        from ...ops import attention as A
        spaces += [(K, _KERNEL_RULES), (A, _ATTENTION_RULES)]
    except Exception:
        pass
    return spaces


def wrapFunc(func, funcFlopCompute):
    """`func` that also charges funcFlopCompute(*args) MACs to the module executing at the time of the call."""
    name = func.__name__

    @wraps(func)
    def counted(*args, **kwds):
        # only the outermost counted call charges: a fused op that falls back to torch.nn.functional on the CPU must
        # not be counted a second time through the functional it calls
        if module_flop_count and _depth[0] == 0:
            module_flop_count[-1].append((name, funcFlopCompute(*args, **kwds)))
        _depth[0] += 1
        try:
            return func(*args, **kwds)
        finally:
            _depth[0] -= 1

    counted.__wrapped_by_flops_profiler__ = func
    return counted


def _patch_functionals():
    for space, rules in _namespaces():
        for name, rule in rules.items():
            fn = getattr(space, name, None)
            if fn is not None and not hasattr(fn, "__wrapped_by_flops_profiler__"):
                old_functions[(space.__name__, name)] = fn
                setattr(space, name, wrapFunc(fn, rule))


def _reload_functionals():
    for space, rules in _namespaces():
        for name in rules:
            original = old_functions.pop((space.__name__, name), None)
            if original is not None:
                setattr(space, name, original)


class _Record:
    """What the profiler knows about one module."""
    __slots__ = ("macs", "params", "seconds", "events", "t0", "handles")

    def __init__(self, module):
        self.macs = 0
        self.params = sum(p.numel() for p in module.parameters() if p.requires_grad)
        self.seconds = 0.0
        self.events = []
        self.t0 = None
        self.handles = []

    def settle(self):
        """Fold finished CUDA event pairs into `seconds`."""
        if self.events:
            torch.cuda.synchronize()
            self.seconds += sum(a.elapsed_time(b) for a, b in self.events) * 1e-3
            self.events = []
        return self.seconds


class FlopsProfiler(object):
    """Measures the latency, estimated number of MACs and parameters of each module in a PyTorch model."""

    def __init__(self, model):
        self.model = model
        self.started = False
        self.func_patched = False
        self.records = {}

    # ---- hooks
    def _attach(self, module, ignore):
        if ignore and type(module) in ignore:
            return
        rec = self.records[module]
        if type(module) in MODULE_HOOK_MAPPING:
            rec.handles.append(module.register_forward_hook(MODULE_HOOK_MAPPING[type(module)]))
            return
        on_gpu = torch.cuda.is_available()

        def enter(mod, inputs):
            module_flop_count.append([])
            if on_gpu:
                rec.t0 = torch.cuda.Event(enable_timing=True)
                rec.t0.record()
            else:
                rec.t0 = time.time()

        def leave(mod, inputs, output):
            if on_gpu:
                end = torch.cuda.Event(enable_timing=True)
                end.record()
                rec.events.append((rec.t0, end))
            else:
                rec.seconds += time.time() - rec.t0
            if module_flop_count:
                rec.macs += sum(m for _, m in module_flop_count.pop())

        rec.handles += [module.register_forward_pre_hook(enter), module.register_forward_hook(leave)]

    def start_profile(self, ignore_list=None):
        self.reset_profile()
        _patch_functionals()
        for module in self.model.modules():
            self._attach(module, ignore_list)
        self.started = self.func_patched = True

    def stop_profile(self):
        if self.started and self.func_patched:
            _reload_functionals()
            self.func_patched = False
        for rec in self.records.values():
            for h in rec.handles:
                h.remove()
            rec.handles = []

    def reset_profile(self):
        self.records = {m: _Record(m) for m in self.model.modules()}

    def end_profile(self):
        if not self.started:
            return
        self.stop_profile()
        self.started = False
        self.records = {}

    # ---- numbers
    def _macs(self, module):
        return self.records[module].macs + sum(self._macs(c) for c in module.children())

    def _seconds(self, module):
        own = self.records[module].settle()
        return own if own != 0 else sum(self.records[c].settle() for c in module.children())

    def get_total_flops(self, as_string=False):
        total = self._macs(self.model)
        return macs_to_string(total) if as_string else total

    def get_total_duration(self, as_string=False):
        total = self.records[self.model].settle()
        return duration_to_string(total) if as_string else total

    def get_total_params(self, as_string=False):
        total = self.records[self.model].params
        return params_to_string(total) if as_string else total

    # ---- report
    def _line(self, module, totals):
        total_macs, total_seconds, total_params = totals
        rec = self.records[module]
        macs, seconds = self._macs(module), self._seconds(module)
        share = lambda part, whole: 0.0 if whole == 0 else part / whole
        return ", ".join([params_to_string(rec.params), "{:.2%} Params".format(share(rec.params, total_params)),
                          macs_to_string(macs), "{:.2%} MACs".format(share(macs, total_macs)),
                          duration_to_string(seconds), "{:.2%} latency".format(share(seconds, total_seconds)),
                          flops_to_string(0.0 if seconds == 0 else 2 * macs / seconds)])

    def print_model_profile(self, profile_step=1, module_depth=-1, top_modules=3, detailed=True):
        totals = (self.get_total_flops(), self.get_total_duration(), self.get_total_params())
        total_macs, total_seconds, total_params = totals
        self.flops, self.params = total_macs, total_params
        row = "{:<60}  {:<8}".format
        print("\n-------------------------- DeepSpeed Flops Profiler --------------------------")
        print(f"Profile Summary at step {profile_step}:")
        print("Notations:\ndata parallel size (dp_size), model paralel size(mp_size),\nnumber of parameters (params), "
              "number of multiply-accumulate operations(MACs),\nnumber of floating point operations (flops), floating "
              "point operations per second (FLOPS),\nfwd latency (forward propagation latency), bwd latency (backward "
              "propagation latency),\nstep (weights update latency), iter latency (sum of fwd, bwd and step latency)\n")
        print(row("params per gpu: ", params_to_string(total_params)))
        print(row("fwd MACs per GPU: ", macs_to_string(total_macs)))
        print(row("fwd latency: ", duration_to_string(total_seconds)))
        if total_seconds > 0:
            print(row("fwd FLOPS per GPU = 2 * fwd MACs per GPU / fwd latency: ",
                      flops_to_string(2 * total_macs / total_seconds)))
        print("\n----------------------------- Aggregated Profile per GPU -----------------------------")
        self.print_model_aggregated_profile(module_depth=module_depth, top_modules=top_modules)
        if detailed:
            print("\n------------------------------ Detailed Profile per GPU ------------------------------")
            print("Each module profile is listed after its name in the following order: \nparams, percentage of total "
                  "params, MACs, percentage of total MACs, fwd latency, percentage of total fwd latency, fwd FLOPS")
            # annotate extra_repr for the duration of one print(model), then put the originals back
            saved = {}
            for m in self.model.modules():
                saved[m] = m.extra_repr
                m.extra_repr = (lambda mod=m, orig=m.extra_repr: ", ".join(
                    s for s in (self._line(mod, totals), orig()) if s))
            try:
                print(self.model)
            finally:
                for m, orig in saved.items():
                    if "extra_repr" in m.__dict__:
                        del m.__dict__["extra_repr"]
        print("------------------------------------------------------------------------------")

    def print_model_aggregated_profile(self, module_depth=-1, top_modules=3):
        if not self.records:
            print("no __flops__ attribute in the model, call this function after start_profile and before end_profile")
            return
        by_depth = {}

        def walk(module, depth):
            slot = by_depth.setdefault(depth, {}).setdefault(module.__class__.__name__, [0, 0, 0.0])
            slot[0] += self._macs(module)
            slot[1] += self.records[module].params
            slot[2] += self._seconds(module)
            for child in module.children():
                walk(child, depth + 1)

        walk(self.model, 0)
        depth = len(by_depth) - 1 if module_depth == -1 else module_depth
        print(f"Top {top_modules} modules in terms of params, MACs or fwd latency at different model depths:")
        for d in range(depth):
            def top(col, fmt):
                ranked = sorted(by_depth[d].items(), key=lambda kv: kv[1][col], reverse=True)[:top_modules]
                return {k: fmt(v[col]) for k, v in ranked}
            print(f"depth {d}:")
            print(f"    params      - {top(1, params_to_string)}")
            print(f"    MACs        - {top(0, macs_to_string)}")
            print(f"    fwd latency - {top(2, duration_to_string)}")


# ----------------------------------------------------------------------------------------------- formatting
def _scaled(value, steps, units, precision, bare_suffix):
    """`value` with the largest unit of `steps` ((factor, suffix), descending) it reaches, or in `units` if given."""
    if units is not None:
        factor = dict((s, f) for f, s in steps).get(units, 1)
        return str(round(value / factor, precision)) + " " + units
    for factor, suffix in steps:
        if value // factor > 0:
            return str(round(value / factor, precision)) + " " + suffix
    return str(value) + bare_suffix


def num_to_string(num, precision=2):
    return _scaled(num, ((10.0 ** 9, "G"), (10.0 ** 6, "M"), (10.0 ** 3, "K")), None, precision, "")


def macs_to_string(macs, units=None, precision=2):
    return _scaled(macs, ((10.0 ** 9, "GMACs"), (10.0 ** 6, "MMACs"), (10.0 ** 3, "KMACs")), units, precision, " MACs")


def number_to_string(num, units=None, precision=2):
    return _scaled(num, ((10.0 ** 9, "G"), (10.0 ** 6, "M"), (10.0 ** 3, "K")), units, precision, " ")


def flops_to_string(flops, units=None, precision=2):
    return _scaled(flops, ((10.0 ** 12, "TFLOPS"), (10.0 ** 9, "GFLOPS"), (10.0 ** 6, "MFLOPS"), (10.0 ** 3, "KFLOPS")),
                   units, precision, " FLOPS")


def params_to_string(params_num, units=None, precision=2):
    if units is not None:
        return str(round(params_num / {"M": 10 ** 6, "K": 10 ** 3}.get(units, 1), precision)) + " " + units
    if params_num // 10 ** 6 > 0:
        return str(round(params_num / 10 ** 6, 2)) + " M"
    if params_num // 10 ** 3:
        return str(round(params_num / 10 ** 3, 2)) + " k"
    return str(params_num)


def duration_to_string(duration, units=None, precision=2):
    if units is not None:
        return str(round(duration * {"us": 10 ** 6, "ms": 10 ** 3}.get(units, 1), precision)) + " " + units
    for factor, suffix in ((1, "s"), (10 ** 3, "ms"), (10 ** 6, "us")):
        if duration * factor > 1:
            return str(round(duration * factor, precision)) + " " + suffix
    return str(duration)


def get_module_flops(module, profiler=None):
    """MACs of `module` and its descendants under `profiler` (the one currently attached, if not given)."""
    prof = profiler or _ACTIVE[-1]
    return prof._macs(module)


def get_module_duration(module, profiler=None):
    prof = profiler or _ACTIVE[-1]
    return prof._seconds(module)


_ACTIVE = []


def get_model_profile(model, input_res, input_constructor=None, print_profile=True, detailed=True, module_depth=-1,
                      top_modules=3, warm_up=1, as_string=True, ignore_modules=None):
    """(MACs, params) of one forward pass on an input of shape `input_res` (batch dimension included)."""
    assert type(input_res) is tuple and len(input_res) >= 1 and isinstance(model, nn.Module)

    def run_once():
        if input_constructor:
            return model(**input_constructor(input_res))
        first = next(model.parameters(), None)
        like = dict(dtype=first.dtype, device=first.device) if first is not None else {}
        return model(torch.empty(tuple(input_res), **like))

    prof = FlopsProfiler(model)
    model.eval()
    for _ in range(warm_up):
        run_once()
    prof.start_profile(ignore_list=ignore_modules)
    _ACTIVE.append(prof)
    try:
        run_once()
        macs, params = prof.get_total_flops(), prof.get_total_params()
        if print_profile:
            prof.print_model_profile(profile_step=warm_up, module_depth=module_depth, top_modules=top_modules,
                                     detailed=detailed)
    finally:
        _ACTIVE.pop()
        prof.end_profile()
    return (macs_to_string(macs), params_to_string(params)) if as_string else (macs, params)
