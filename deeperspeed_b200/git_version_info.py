"""Version info module of the reference (deepspeed/git_version_info.py:1-17): version string, git hash / branch and
the op compatibility tables that `ds_report` prints."""
from .version import __version__ as version

git_hash = "b200-native"
git_branch = "main"
try:
    from .ops.op_builder import ALL_OPS
    installed_ops = {name: True for name in ALL_OPS}          # everything is built ahead of time, in tree
    compatible_ops = {name: b.is_compatible() for name, b in ALL_OPS.items()}
except Exception:                                             # pragma: no cover
    installed_ops, compatible_ops = {}, {}
