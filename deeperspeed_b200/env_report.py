"""`ds_report`: which native ops are built / compatible, plus torch / CUDA / nvcc versions. Same two tables and the
same entry points as the reference tool (deepspeed/env_report.py); the rows describe THIS build: one ahead-of-time
sm_100a library for the CUDA ops, one host library for the C++ ops."""
import os
import re
import shutil
import subprocess

import torch

from .ops import native as N
from .ops.op_builder import ALL_OPS
from .version import __version__, git_hash, git_branch

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
color_len = len(GREEN) + len(END)


def _tag(colour, text):
    return f"{colour}{text}{END}"


SUCCESS, WARNING = _tag(GREEN, " [SUCCESS] "), _tag(YELLOW, " [WARNING] ")
OKAY, FAIL, INFO = _tag(GREEN, "[OKAY]"), _tag(RED, "[FAIL]"), "[INFO]"
okay, warning = OKAY, _tag(YELLOW, "[WARNING]")
_YES, _NO = _tag(GREEN, "[YES]"), _tag(YELLOW, "[NO]")


def _visible_len(cell):
    return len(re.sub(r"\033\[\d+m", "", cell))


def _row(cells, widths):
    """Cells joined by dot leaders so that column i starts `widths[i]` visible characters after column i - 1."""
    out = []
    for cell, width in zip(cells, widths):
        out.append(cell + " " + "." * max(width - _visible_len(cell) - 1, 1) + " ")
    return "".join(out) + cells[-1] if len(cells) > len(widths) else "".join(out).rstrip(" .")


def op_report():
    widths = (24, 12)
    rule = "-" * (sum(widths) + len("compatible"))
    built = {True: os.path.exists(N.lib_path()), False: os.path.exists(N.lib_path("libdsb200_host.so"))}
    print(rule)
    print("DeepSpeed C++/CUDA extension op report (B200-native build: sm_100a only, ahead of time)")
    print(rule)
    print(_row(["op name", "installed", "compatible"], widths))
    print(rule)
    for name, builder in ALL_OPS.items():
        print(_row([name, _YES if built[bool(builder.CUDA)] else _NO, OKAY if builder.is_compatible() else _NO], widths))
    print(rule)


def nvcc_version():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    try:
        banner = subprocess.check_output([nvcc, "-V"], universal_newlines=True)
    except Exception:
        return _tag(RED, " [FAIL] nvcc missing ")
    found = re.search(r"release\s+([0-9.]+)", banner)
    return found.group(1) if found else "unknown"


def debug_report():
    cuda_lib = N.lib_path()
    facts = {
        "torch install path": torch.__path__,
        "torch version": torch.__version__,
        "torch cuda version": torch.version.cuda,
        "nvcc version": nvcc_version(),
        "deepspeed info": f"{__version__}, {git_hash}, {git_branch}",
        "target architecture": "sm_100a (compute_100a)",
        "cuda library": cuda_lib if os.path.exists(cuda_lib) else "not built",
        "gpu": torch.cuda.get_device_name(0) if torch.cuda.is_available() else "none visible",
    }
    print("DeepSpeed general environment info:")
    for name, value in facts.items():
        print(_row([name, str(value)], (34,)))


def ninja_installed():
    try:
        import ninja  # noqa: F401
    except ImportError:
        return False
    return True


def main():
    op_report()
    debug_report()


if __name__ == "__main__":
    main()
