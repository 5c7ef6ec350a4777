from . import adam, lamb

from ..git_version_info import compatible_ops as __compatible_ops__  # noqa: E402,F401
