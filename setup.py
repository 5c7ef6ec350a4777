"""Packaging for deeperspeed_b200 (counterpart of the reference's setup.py / install.sh).

`pip install .` compiles the native libraries ahead of time for sm_100a (tools/build_native.py: nvcc for the CUDA
kernels, g++ for the host SIMD / AIO code) and ships them inside the package; there is no JIT path and no other
architecture. DSB_SKIP_NATIVE=1 skips the build (CPU-only checkouts: the pure-Python engine still works, GPU ops raise).
"""
import os
import subprocess
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


def read_version():
    scope = {}
    with open(os.path.join(HERE, "deeperspeed_b200", "version.py")) as f:
        exec(f.read(), scope)
    return scope.get("__version__", "0.0.0")


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("DSB_SKIP_NATIVE", "0") != "1":
            subprocess.check_call([sys.executable, os.path.join(HERE, "tools", "build_native.py")], cwd=HERE)
        super().run()


def requirements(name):
    path = os.path.join(HERE, "requirements", name)
    if not os.path.exists(path):
        return []
    return [line.strip() for line in open(path) if line.strip() and not line.startswith("#")]


setup(
    name="deeperspeed_b200",
    version=read_version(),
    description="B200-native (sm_100a) DeepSpeed-compatible training engine",
    packages=find_packages(include=["deeperspeed_b200", "deeperspeed_b200.*"]),
    package_data={"deeperspeed_b200": ["lib/*.so"]},
    include_package_data=True,
    scripts=["bin/deepspeed", "bin/deepspeed.pt", "bin/ds", "bin/ds_ssh", "bin/ds_report", "bin/ds_elastic"],
    install_requires=requirements("requirements.txt"),
    extras_require={"dev": requirements("requirements-dev.txt"), "1bit_mpi": requirements("requirements-1bit-mpi.txt")},
    python_requires=">=3.10",
    cmdclass={"build_py": BuildWithNative},
)
