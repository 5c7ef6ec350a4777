"""Run the REFERENCE's own unit-test files, unmodified, against this package.

    python tools/run_reference_unit_tests.py [--ref /root/reference] [--files a.py,b.py] [pytest args ...]

`import deepspeed` inside those files resolves to deeperspeed_b200 (`install_deepspeed_alias`). The reference's fixtures
(`simple_model.py`, `multi_output_model.py`, the JSON configs) are used as they are; only its process harness
`common.distributed_test` - which wants NCCL and one GPU per rank - is replaced by this repo's `tests/common.py`, so
the files also run on a CPU-only machine (gloo). The working directory is the reference checkout because some tests
open `tests/unit/ds_batch_config.json` by relative path; nothing is written there (`-p no:cacheprovider`).

This is a drop-in check, not part of `pytest tests/`: it needs the reference tree. Results of the last run are
recorded in `profiles/reference_unit_tests.txt`."""
import argparse
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# CPU-runnable files (the kernel files - test_cuda_forward/backward, test_sparse_attention, test_cpu_adam, test_onebit,
# test_adamw - build CUDA / Triton ops of the reference or need GPUs; their role is covered by tests/*_gpu.py)
DEFAULT_FILES = ["test_config.py", "test_lr_schedulers.py", "test_dynamic_loss_scale.py", "test_pld.py",
                 "test_multi_output_model.py", "test_zero.py", "test_zero_context.py", "test_data.py",
                 "test_flops_profiler.py", "test_activation_checkpointing.py", "test_pipe_module.py", "test_topology.py",
                 "test_partition.py", "test_dist.py", "test_pipe_schedule.py", "test_elastic.py", "test_run.py",
                 "test_ds_arguments.py", "test_zero_tiled.py", "test_runtime_utils.py", "test_csr.py", "test_fp16.py",
                 "test_checkpointing.py"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default=os.environ.get("DSB_REFERENCE", "/root/reference"))
    ap.add_argument("--files", default=",".join(DEFAULT_FILES))
    args, pytest_args = ap.parse_known_args()
    unit = os.path.join(args.ref, "tests", "unit")
    assert os.path.isdir(unit), f"no reference unit tests under {unit}"

    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.join(HERE, "tests"))
    import deeperspeed_b200
    deeperspeed_b200.install_deepspeed_alias()
    import common                                   # this repo's gloo / CPU harness ...
    sys.modules["common"] = common                  # ... under the name the reference's files import
    for name in ("simple_model", "multi_output_model"):      # the reference's fixtures, not ours
        spec = importlib.util.spec_from_file_location(name, os.path.join(unit, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
    import torch
    if not torch.cuda.is_available():
        # the reference's fixture asserts world_size <= torch.cuda.device_count() (one GPU per rank); on a CPU-only
        # machine the ranks are gloo processes, so that count is meaningless - let the assertion pass
        torch.cuda.device_count = lambda: 64
        # ... and `.cuda()` / `.to("cuda")` in the test bodies keep the tensor / module where it is, so the logic
        # under test (activation checkpointing, PartitionedTensor, grids, pipeline modules) runs on CPU tensors
        def _is_cuda_target(args, kwargs):
            dev = kwargs.get("device", args[0] if args else None)
            return isinstance(dev, (str, torch.device)) and str(dev).startswith("cuda")

        tensor_to, module_to = torch.Tensor.to, torch.nn.Module.to
        torch.Tensor.cuda = lambda self, *a, **k: self
        torch.nn.Module.cuda = lambda self, *a, **k: self
        torch.Tensor.to = lambda self, *a, **k: self if _is_cuda_target(a, k) else tensor_to(self, *a, **k)
        torch.nn.Module.to = lambda self, *a, **k: self if _is_cuda_target(a, k) else module_to(self, *a, **k)
    os.chdir(args.ref)
    import inspect
    import pytest

    class InProcessInitLast:
        """Tests that initialise torch.distributed in the pytest process itself (`setup_serial_env()` in the
        reference's files) go last: the reference runs its suite with `--forked`, here the fork-based tests must
        not inherit that group."""

        @staticmethod
        def pytest_collection_modifyitems(items):
            def in_process(item):
                try:
                    return "setup_serial_env" in inspect.getsource(item.function)
                except (OSError, TypeError, AttributeError):
                    return False
            items.sort(key=in_process)          # stable: False (fork-based) before True

    files = [os.path.join(unit, f) for f in args.files.split(",") if f]
    return pytest.main(["-p", "no:cacheprovider", "-q", "--no-header", "--rootdir=/tmp"] + pytest_args + files,
                       plugins=[InProcessInitLast()])


if __name__ == "__main__":
    sys.exit(main())
