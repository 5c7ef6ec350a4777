"""Version info module of the reference (deepspeed/git_version_info.py:1-17): version string, git hash / branch and
the op compatibility tables that `ds_report` prints."""
import torch

from .version import __version__ as version, git_hash, git_branch

torch_info = {"version": ".".join(torch.__version__.split(".")[:2]), "cuda_version": torch.version.cuda or "0.0"}
try:
    from .ops.op_builder import ALL_OPS
    installed_ops = {name: True for name in ALL_OPS}          # everything is built ahead of time, in tree
    compatible_ops = {name: b.is_compatible() for name, b in ALL_OPS.items()}
except Exception:                                             # pragma: no cover
    installed_ops, compatible_ops = {}, {}
