"""FlopsProfiler: counts MACs/params/latency per module by wrapping torch.nn.functional ops and module hooks; prints the
annotated module tree and depth-aggregated top-k (API of reference deepspeed/profiling/flops_profiler/profiler.py:11-868)."""
import time
from functools import partial

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

module_flop_count = []
old_functions = {}


def _prod(dims):
    p = 1
    for v in dims:
        p *= v
    return p


def _linear_flops_compute(input, weight, bias=None):
    return _prod(input.shape) * weight.shape[0]


def _relu_flops_compute(input, inplace=False):
    return input.numel()


def _elementwise(input, *args, **kwargs):
    return input.numel()


def _pool_flops_compute(input, *args, **kwargs):
    return input.numel()


def _conv_flops_compute(input, weight, bias=None, stride=1, padding=0, dilation=1, groups=1):
    batch_size = input.shape[0]
    in_channels = input.shape[1]
    out_channels = weight.shape[0]
    kernel_dims = list(weight.shape[2:])
    input_dims = list(input.shape[2:])
    length = len(input_dims)
    paddings = padding if type(padding) is tuple else (padding,) * length
    strides = stride if type(stride) is tuple else (stride,) * length
    dilations = dilation if type(dilation) is tuple else (dilation,) * length
    output_dims = []
    for idx, input_dim in enumerate(input_dims):
        pad = paddings[idx] if not isinstance(paddings[idx], str) else 0
        output_dim = (input_dim + 2 * pad - (dilations[idx] * (kernel_dims[idx] - 1) + 1)) // strides[idx] + 1
        output_dims.append(output_dim)
    filters_per_channel = out_channels // groups
    conv_per_position_flops = int(_prod(kernel_dims)) * in_channels * filters_per_channel
    active_elements_count = batch_size * int(_prod(output_dims))
    overall_conv_flops = conv_per_position_flops * active_elements_count
    bias_flops = out_channels * active_elements_count if bias is not None else 0
    return int(overall_conv_flops + bias_flops)


def _norm_flops(input, *args, **kwargs):
    return input.numel() * 5


def _softmax_flops_compute(input, dim=None, _stacklevel=3, dtype=None):
    return input.numel()


def _embedding_flops_compute(input, weight, *args, **kwargs):
    return 0


def _matmul_flops(input, other, *args, **kwargs):
    return _prod(input.shape) * other.shape[-1]


def _sdpa_flops(q, k, v, *args, **kwargs):
    return 2 * _prod(q.shape) * k.shape[-2]


MODULE_HOOK_MAPPING = {}


def wrapFunc(func, funcFlopCompute):
    oldFunc = func
    name = func.__name__
    old_functions[name] = oldFunc

    def newFunc(*args, **kwds):
        flops = funcFlopCompute(*args, **kwds)
        if module_flop_count:
            module_flop_count[-1].append((name, flops))
        return oldFunc(*args, **kwds)

    newFunc.__name__ = name
    return newFunc


_F_PATCHES = {
    "linear": _linear_flops_compute, "conv1d": _conv_flops_compute, "conv2d": _conv_flops_compute,
    "conv3d": _conv_flops_compute, "relu": _relu_flops_compute, "gelu": _elementwise, "silu": _elementwise,
    "elu": _elementwise, "leaky_relu": _elementwise, "relu6": _elementwise, "prelu": _elementwise,
    "batch_norm": _norm_flops, "layer_norm": _norm_flops, "instance_norm": _norm_flops, "group_norm": _norm_flops,
    "avg_pool1d": _pool_flops_compute, "avg_pool2d": _pool_flops_compute, "avg_pool3d": _pool_flops_compute,
    "max_pool1d": _pool_flops_compute, "max_pool2d": _pool_flops_compute, "max_pool3d": _pool_flops_compute,
    "adaptive_avg_pool2d": _pool_flops_compute, "adaptive_max_pool2d": _pool_flops_compute,
    "softmax": _softmax_flops_compute, "embedding": _embedding_flops_compute,
    "scaled_dot_product_attention": _sdpa_flops,
}


def _patch_functionals():
    for name, fn in _F_PATCHES.items():
        if hasattr(F, name):
            setattr(F, name, wrapFunc(getattr(F, name), fn))


def _reload_functionals():
    for name in _F_PATCHES:
        if name in old_functions:
            setattr(F, name, old_functions[name])


class FlopsProfiler(object):
    """Measures the latency, estimated number of MACs and parameters of each module in a PyTorch model."""

    def __init__(self, model):
        self.model = model
        self.started = False
        self.func_patched = False

    def start_profile(self, ignore_list=None):
        self.reset_profile()
        _patch_functionals()

        def register_module_hooks(module, ignore_list):
            if ignore_list and type(module) in ignore_list:
                return
            if type(module) in MODULE_HOOK_MAPPING:
                module.__flops_handle__ = module.register_forward_hook(MODULE_HOOK_MAPPING[type(module)])
                return

            def pre_hook(module, input):
                module_flop_count.append([])

            module.__pre_hook_handle__ = module.register_forward_pre_hook(pre_hook)

            def post_hook(module, input, output):
                if module_flop_count:
                    module.__flops__ += sum([elem[1] for elem in module_flop_count[-1]])
                    module_flop_count.pop()

            module.__post_hook_handle__ = module.register_forward_hook(post_hook)

            def start_time_hook(module, input):
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                module.__start_time__ = time.time()

            module.__start_time_hook_handle__ = module.register_forward_pre_hook(start_time_hook)

            def end_time_hook(module, input, output):
                if torch.cuda.is_available():
                    torch.cuda.synchronize()
                module.__duration__ += time.time() - module.__start_time__

            module.__end_time_hook_handle__ = module.register_forward_hook(end_time_hook)

        self.model.apply(partial(register_module_hooks, ignore_list=ignore_list))
        self.started = True
        self.func_patched = True

    def stop_profile(self):
        if self.started and self.func_patched:
            _reload_functionals()
            self.func_patched = False

        def remove_profile_attrs(module):
            for attr in ("__pre_hook_handle__", "__post_hook_handle__", "__flops_handle__",
                         "__start_time_hook_handle__", "__end_time_hook_handle__"):
                if hasattr(module, attr):
                    getattr(module, attr).remove()
                    delattr(module, attr)

        self.model.apply(remove_profile_attrs)

    def reset_profile(self):
        def add_or_reset_attrs(module):
            module.__flops__ = 0
            module.__params__ = sum(p.numel() for p in module.parameters() if p.requires_grad)
            module.__start_time__ = 0
            module.__duration__ = 0

        self.model.apply(add_or_reset_attrs)

    def end_profile(self):
        if not self.started:
            return
        self.stop_profile()
        self.started = False

        def remove_profile_attrs(module):
            for attr in ("__flops__", "__params__", "__start_time__", "__duration__"):
                if hasattr(module, attr):
                    delattr(module, attr)

        self.model.apply(remove_profile_attrs)

    def get_total_flops(self, as_string=False):
        total_flops = get_module_flops(self.model)
        return macs_to_string(total_flops) if as_string else total_flops

    def get_total_duration(self, as_string=False):
        total_duration = self.model.__duration__
        return duration_to_string(total_duration) if as_string else total_duration

    def get_total_params(self, as_string=False):
        return params_to_string(self.model.__params__) if as_string else self.model.__params__

    def print_model_profile(self, profile_step=1, module_depth=-1, top_modules=3, detailed=True):
        total_flops = self.get_total_flops()
        total_duration = self.get_total_duration()
        total_params = self.get_total_params()
        self.flops = total_flops
        self.params = total_params
        print("\n-------------------------- DeepSpeed Flops Profiler --------------------------")
        print(f"Profile Summary at step {profile_step}:")
        print("Notations:\ndata parallel size (dp_size), model paralel size(mp_size),\nnumber of parameters (params), "
              "number of multiply-accumulate operations(MACs),\nnumber of floating point operations (flops), floating "
              "point operations per second (FLOPS),\nfwd latency (forward propagation latency), bwd latency (backward "
              "propagation latency),\nstep (weights update latency), iter latency (sum of fwd, bwd and step latency)\n")
        print("{:<60}  {:<8}".format("params per gpu: ", params_to_string(total_params)))
        print("{:<60}  {:<8}".format("fwd MACs per GPU: ", macs_to_string(total_flops)))
        print("{:<60}  {:<8}".format("fwd latency: ", duration_to_string(total_duration)))
        if total_duration > 0:
            print("{:<60}  {:<8}".format("fwd FLOPS per GPU = 2 * fwd MACs per GPU / fwd latency: ",
                                         flops_to_string(2 * total_flops / total_duration)))

        def flops_repr(module):
            params = module.__params__
            flops = get_module_flops(module)
            items = [params_to_string(params), "{:.2%} Params".format(params / total_params if total_params else 0),
                     macs_to_string(flops), "{:.2%} MACs".format(0.0 if total_flops == 0 else flops / total_flops)]
            duration = module.__duration__
            if duration == 0:
                for m in module.children():
                    duration += m.__duration__
            items.append(duration_to_string(duration))
            items.append("{:.2%} latency".format(0.0 if total_duration == 0 else duration / total_duration))
            items.append(flops_to_string(0.0 if duration == 0 else 2 * flops / duration))
            items.append(module.original_extra_repr())
            return ", ".join(items)

        def add_extra_repr(module):
            flops_extra_repr = flops_repr.__get__(module)
            if module.extra_repr != flops_extra_repr:
                module.original_extra_repr = module.extra_repr
                module.extra_repr = flops_extra_repr
                assert module.extra_repr != module.original_extra_repr

        def del_extra_repr(module):
            if hasattr(module, "original_extra_repr"):
                module.extra_repr = module.original_extra_repr
                del module.original_extra_repr

        self.model.apply(add_extra_repr)
        print("\n----------------------------- Aggregated Profile per GPU -----------------------------")
        self.print_model_aggregated_profile(module_depth=module_depth, top_modules=top_modules)
        if detailed:
            print("\n------------------------------ Detailed Profile per GPU ------------------------------")
            print("Each module profile is listed after its name in the following order: \nparams, percentage of total "
                  "params, MACs, percentage of total MACs, fwd latency, percentage of total fwd latency, fwd FLOPS")
            print(self.model)
        self.model.apply(del_extra_repr)
        print("------------------------------------------------------------------------------")

    def print_model_aggregated_profile(self, module_depth=-1, top_modules=3):
        info = {}
        if not hasattr(self.model, "__flops__"):
            print("no __flops__ attribute in the model, call this function after start_profile and before end_profile")
            return

        def walk_module(module, curr_depth, info):
            if curr_depth not in info:
                info[curr_depth] = {}
            if module.__class__.__name__ not in info[curr_depth]:
                info[curr_depth][module.__class__.__name__] = [0, 0, 0]
            info[curr_depth][module.__class__.__name__][0] += get_module_flops(module)
            info[curr_depth][module.__class__.__name__][1] += module.__params__
            info[curr_depth][module.__class__.__name__][2] += get_module_duration(module)
            has_children = len(module._modules.items()) != 0
            if has_children:
                for child in module.children():
                    walk_module(child, curr_depth + 1, info)

        walk_module(self.model, 0, info)
        depth = module_depth
        if module_depth == -1:
            depth = len(info) - 1
        print(f"Top {top_modules} modules in terms of params, MACs or fwd latency at different model depths:")
        for d in range(depth):
            num_items = min(top_modules, len(info[d]))
            sort_macs = {k: macs_to_string(v[0]) for k, v in sorted(info[d].items(), key=lambda item: item[1][0], reverse=True)[:num_items]}
            sort_params = {k: params_to_string(v[1]) for k, v in sorted(info[d].items(), key=lambda item: item[1][1], reverse=True)[:num_items]}
            sort_time = {k: duration_to_string(v[2]) for k, v in sorted(info[d].items(), key=lambda item: item[1][2], reverse=True)[:num_items]}
            print(f"depth {d}:")
            print(f"    params      - {sort_params}")
            print(f"    MACs        - {sort_macs}")
            print(f"    fwd latency - {sort_time}")


def num_to_string(num, precision=2):
    if num // 10 ** 9 > 0:
        return str(round(num / 10.0 ** 9, precision)) + " G"
    elif num // 10 ** 6 > 0:
        return str(round(num / 10.0 ** 6, precision)) + " M"
    elif num // 10 ** 3 > 0:
        return str(round(num / 10.0 ** 3, precision)) + " K"
    return str(num)


def macs_to_string(macs, units=None, precision=2):
    if units is None:
        if macs // 10 ** 9 > 0:
            return str(round(macs / 10.0 ** 9, precision)) + " GMACs"
        elif macs // 10 ** 6 > 0:
            return str(round(macs / 10.0 ** 6, precision)) + " MMACs"
        elif macs // 10 ** 3 > 0:
            return str(round(macs / 10.0 ** 3, precision)) + " KMACs"
        return str(macs) + " MACs"
    scale = {"GMACs": 10.0 ** 9, "MMACs": 10.0 ** 6, "KMACs": 10.0 ** 3}.get(units, 1)
    return str(round(macs / scale, precision)) + " " + units


def number_to_string(num, units=None, precision=2):
    if units is None:
        if num // 10 ** 9 > 0:
            return str(round(num / 10.0 ** 9, precision)) + " G"
        elif num // 10 ** 6 > 0:
            return str(round(num / 10.0 ** 6, precision)) + " M"
        elif num // 10 ** 3 > 0:
            return str(round(num / 10.0 ** 3, precision)) + " K"
        return str(num) + " "
    scale = {"G": 10.0 ** 9, "M": 10.0 ** 6, "K": 10.0 ** 3}.get(units, 1)
    return str(round(num / scale, precision)) + " " + units


def flops_to_string(flops, units=None, precision=2):
    if units is None:
        if flops // 10 ** 12 > 0:
            return str(round(flops / 10.0 ** 12, precision)) + " TFLOPS"
        if flops // 10 ** 9 > 0:
            return str(round(flops / 10.0 ** 9, precision)) + " GFLOPS"
        elif flops // 10 ** 6 > 0:
            return str(round(flops / 10.0 ** 6, precision)) + " MFLOPS"
        elif flops // 10 ** 3 > 0:
            return str(round(flops / 10.0 ** 3, precision)) + " KFLOPS"
        return str(flops) + " FLOPS"
    scale = {"TFLOPS": 10.0 ** 12, "GFLOPS": 10.0 ** 9, "MFLOPS": 10.0 ** 6, "KFLOPS": 10.0 ** 3}.get(units, 1)
    return str(round(flops / scale, precision)) + " " + units


def params_to_string(params_num, units=None, precision=2):
    if units is None:
        if params_num // 10 ** 6 > 0:
            return str(round(params_num / 10 ** 6, 2)) + " M"
        elif params_num // 10 ** 3:
            return str(round(params_num / 10 ** 3, 2)) + " k"
        return str(params_num)
    scale = {"M": 10 ** 6, "K": 10 ** 3}.get(units, 1)
    return str(round(params_num / scale, precision)) + " " + units


def duration_to_string(duration, units=None, precision=2):
    if units is None:
        if duration > 1:
            return str(round(duration, precision)) + " s"
        elif duration * 10 ** 3 > 1:
            return str(round(duration * 10 ** 3, precision)) + " ms"
        elif duration * 10 ** 6 > 1:
            return str(round(duration * 10 ** 6, precision)) + " us"
        return str(duration)
    scale = {"us": 10 ** 6, "ms": 10 ** 3}.get(units, 1)
    return str(round(duration * scale, precision)) + " " + units


def get_module_flops(module):
    total = module.__flops__
    for child in module.children():
        total += get_module_flops(child)
    return total


def get_module_duration(module):
    duration = module.__duration__
    if duration == 0:
        for m in module.children():
            duration += m.__duration__
    return duration


def get_model_profile(model, input_res, input_constructor=None, print_profile=True, detailed=True, module_depth=-1,
                      top_modules=3, warm_up=1, as_string=True, ignore_modules=None):
    """Returns (MACs, params) of one forward pass on an input of shape `input_res` (batch included)."""
    assert type(input_res) is tuple and len(input_res) >= 1 and isinstance(model, nn.Module)
    prof = FlopsProfiler(model)
    model.eval()
    for _ in range(warm_up):
        if input_constructor:
            _ = model(**input_constructor(input_res))
        else:
            try:
                batch = torch.ones(()).new_empty((*input_res,), dtype=next(model.parameters()).dtype,
                                                 device=next(model.parameters()).device)
            except StopIteration:
                batch = torch.ones(()).new_empty((*input_res,))
            _ = model(batch)
    prof.start_profile(ignore_list=ignore_modules)
    if input_constructor:
        _ = model(**input_constructor(input_res))
    else:
        try:
            batch = torch.ones(()).new_empty((*input_res,), dtype=next(model.parameters()).dtype,
                                             device=next(model.parameters()).device)
        except StopIteration:
            batch = torch.ones(()).new_empty((*input_res,))
        _ = model(batch)
    flops = prof.get_total_flops()
    params = prof.get_total_params()
    if print_profile:
        prof.print_model_profile(profile_step=warm_up, module_depth=module_depth, top_modules=top_modules, detailed=detailed)
    prof.end_profile()
    if as_string:
        return macs_to_string(flops), params_to_string(params)
    return flops, params
