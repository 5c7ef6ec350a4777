#!/usr/bin/env python
"""Ahead-of-time build of the native libraries (no JIT cache: the .so files live in-tree and ship with gpurun).

  libdsb200.so       all CUDA kernels, compiled for sm_100a only (nvcc -gencode arch=compute_100a,code=sm_100a)
  libdsb200_host.so  host runtime pieces (CPU Adam, async I/O, LUT builders, flatten) with g++ -O3 -fopenmp

Covers what the reference does with op_builder JIT/prebuilt extension modules (reference op_builder/builder.py:81-340)
but as one deterministic AOT step.
"""
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
OUT = os.path.join(ROOT, "deeperspeed_b200", "lib")
OBJ = os.path.join(ROOT, "build", "obj")

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
CUDA_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "--use_fast_math",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]
HOST_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-Wall", "-pthread"]


def _sources(sub_exts):
    out = []
    for dirpath, _, files in os.walk(CSRC):
        for f in sorted(files):
            if f.endswith(sub_exts):
                out.append(os.path.join(dirpath, f))
    return sorted(out)


def _stale(src, obj, deps):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src] + deps)


def _compile(cmd, log):
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("compile failed: %s\n%s" % (" ".join(cmd), (r.stdout + r.stderr)[-6000:]))
    return r.stderr


def build(verbose=True, force=False):
    os.makedirs(OUT, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    headers = _sources((".cuh", ".h", ".hpp")) + [os.path.join(ROOT, "include", "dsb200.h")]
    cu = _sources((".cu",))
    cpp = [s for s in _sources((".cpp",))]
    jobs = []
    cu_objs, cpp_objs = [], []
    for src in cu:
        obj = os.path.join(OBJ, hashlib.md5(src.encode()).hexdigest()[:8] + "_" + os.path.basename(src) + ".o")
        cu_objs.append(obj)
        if force or _stale(src, obj, headers):
            jobs.append((NVCC, CUDA_FLAGS + ["-c", src, "-o", obj], obj))
    for src in cpp:
        obj = os.path.join(OBJ, hashlib.md5(src.encode()).hexdigest()[:8] + "_" + os.path.basename(src) + ".o")
        cpp_objs.append(obj)
        if force or _stale(src, obj, headers):
            jobs.append(("g++", HOST_FLAGS + ["-c", src, "-o", obj], obj))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
            futs = {ex.submit(_compile, [c] + a, o + ".log"): o for c, a, o in jobs}
            for fut in cf.as_completed(futs):
                fut.result()
                if verbose:
                    print("[build] compiled", os.path.basename(futs[fut]))
    lib_cuda = os.path.join(OUT, "libdsb200.so")
    lib_host = os.path.join(OUT, "libdsb200_host.so")
    if cu_objs and (jobs or not os.path.exists(lib_cuda)):
        _compile([NVCC, "-shared", "-o", lib_cuda] + cu_objs + ["-lcudart", "-Xcompiler", "-fPIC"], lib_cuda + ".log")
    if cpp_objs and (jobs or not os.path.exists(lib_host)):
        _compile(["g++", "-shared", "-o", lib_host] + cpp_objs + ["-fopenmp", "-pthread"], lib_host + ".log")
    if verbose:
        print("[build] ok:", lib_cuda, lib_host if cpp_objs else "")
    return lib_cuda, lib_host


if __name__ == "__main__":
    build(force="--force" in sys.argv)
