"""`ds_report`: which native ops are built / compatible, plus torch / CUDA / nvcc versions
(reference deepspeed/env_report.py:23-109)."""
import os
import shutil
import subprocess

import torch

from .ops import native as N
from .ops.op_builder import ALL_OPS
from .version import __version__, git_hash, git_branch

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
SUCCESS = f"{GREEN} [SUCCESS] {END}"
OKAY = f"{GREEN}[OKAY]{END}"
WARNING = f"{YELLOW} [WARNING] {END}"
FAIL = f"{RED}[FAIL]{END}"
INFO = "[INFO]"
color_len = len(GREEN) + len(END)
okay = f"{GREEN}[OKAY]{END}"
warning = f"{YELLOW}[WARNING]{END}"


def op_report():
    max_dots = 23
    max_dots2 = 11
    h = ["op name", "installed", "compatible"]
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    print("DeepSpeed C++/CUDA extension op report (B200-native build: sm_100a only, ahead of time)")
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    print(h[0], "." * (max_dots - len(h[0])), h[1], "." * (max_dots2 - len(h[1])), h[2])
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))
    installed = f"{GREEN}[YES]{END}"
    no = f"{YELLOW}[NO]{END}"
    have_cuda = os.path.exists(N.lib_path())
    have_host = os.path.exists(N.lib_path("libdsb200_host.so"))
    for op_name, builder in ALL_OPS.items():
        dots = "." * (max_dots - len(op_name))
        is_compatible = OKAY if builder.is_compatible() else no
        is_installed = installed if ((have_cuda if builder.CUDA else have_host)) else no
        dots2 = "." * ((len(h[1]) + (max_dots2 - len(h[1]))) - (len(is_installed) - color_len))
        print(op_name, dots, is_installed, dots2, is_compatible)
    print("-" * (max_dots + max_dots2 + len(h[0]) + len(h[1])))


def nvcc_version():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    try:
        output = subprocess.check_output([nvcc, "-V"], universal_newlines=True)
    except Exception:
        return f"{RED} [FAIL] nvcc missing {END}"
    output_split = output.split()
    release_idx = output_split.index("release")
    release = output_split[release_idx + 1].replace(",", "").split(".")
    return ".".join(release)


def debug_report():
    max_dots = 33
    report = [("torch install path", torch.__path__), ("torch version", torch.__version__),
              ("torch cuda version", torch.version.cuda), ("nvcc version", nvcc_version()),
              ("deepspeed info", f"{__version__}, {git_hash}, {git_branch}"),
              ("target architecture", "sm_100a (compute_100a)"),
              ("cuda library", N.lib_path() if os.path.exists(N.lib_path()) else "not built"),
              ("gpu", torch.cuda.get_device_name(0) if torch.cuda.is_available() else "none visible")]
    print("DeepSpeed general environment info:")
    for name, value in report:
        print(name, "." * (max_dots - len(name)), value)


def main():
    op_report()
    debug_report()


if __name__ == "__main__":
    main()


def ninja_installed():
    try:
        import ninja  # noqa: F401
    except ImportError:
        return False
    return True
