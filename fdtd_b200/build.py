"""Builds fdtd_b200/libfdtd_b200.so (CUDA, sm_100a) in-tree with nvcc.  No torch involved."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_DIR, "libfdtd_b200.so")
_SOURCES = [os.path.join(_DIR, "csrc", "fdtd_b200.cu")]
_DEPS = _SOURCES + sorted(glob.glob(os.path.join(_DIR, "csrc", "*.cuh"))) + [os.path.join(os.path.dirname(_DIR), "include", "fdtd_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared",
    "-Xptxas", "-v",
    "--resource-usage",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: fdtd_b200 needs the CUDA toolkit to build its sm_100a library")


def is_fresh() -> bool:
    return os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(p) for p in _DEPS)


def build_extension(force: bool = False, verbose: bool = False) -> str:
    """Compile the shared library if it is missing or older than its sources.  Returns its path."""
    if is_fresh() and not force:
        return LIB_PATH
    # FDTD_B200_NVCC_EXTRA: extra flags for diagnostic builds (e.g. -DFDTD_B200_CTA_TRACE, tools/r02_cta_trace.sh); never set for the shipped library
    cmd = [_nvcc(), *NVCC_FLAGS, *os.environ.get("FDTD_B200_NVCC_EXTRA", "").split(), "-o", LIB_PATH, *_SOURCES, "-lcudart"]
    env = dict(os.environ)
    # the image exports CXX/CC wrappers that lack OpenMP specs; nvcc only needs a plain host g++
    if os.path.exists("/usr/bin/g++"):
        cmd[1:1] = ["-ccbin", "/usr/bin/g++"]
    res = subprocess.run(cmd, capture_output=True, text=True, env=env)
    log = res.stdout + res.stderr
    with open(os.path.join(_DIR, "csrc", "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if verbose:
        print(log)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + log[-4000:])
    return LIB_PATH


if __name__ == "__main__":
    print(build_extension(force=True, verbose=True))
