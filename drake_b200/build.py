"""Builds drake_b200/lib/libgmpm_b200.so (hand-written CUDA for sm_100a + the C ABI) in-tree with nvcc.

`python -m drake_b200.build [--force]`.  nvcc cross-compiles without a GPU.  The built library is
git-ignored but travels to the GPU box with the repository snapshot.
"""
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libgmpm_b200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found (needed to build libgmpm_b200.so)")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh"))
                  + glob.glob(os.path.join(ROOT, "include", "*.h")))


def up_to_date():
    return os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(s) for s in sources())


def build(force=False, verbose=False):
    if not force and up_to_date():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    cmd = [_nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared",
           "-I" + os.path.join(ROOT, "include"), os.path.join(CSRC, "gmpm_capi.cu"), "-o", LIB]
    cmd[1:1] = os.environ.get("GMPM_NVCC_FLAGS", "").split()   # development A/B switches (-DGMPM_CHUNK=120 ...)
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    env = dict(os.environ)
    # the image's CC/CXX wrappers are not usable as nvcc host compilers; let nvcc pick the distro g++
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
