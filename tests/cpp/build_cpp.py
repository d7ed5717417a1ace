"""Builds tests/cpp/_build/test_shim (g++, C++17) against include/ and the two in-tree libraries."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "_build", "test_shim")


def build(force=False):
    src = os.path.join(HERE, "test_shim.cc")
    deps = [src, os.path.join(ROOT, "drake_b200", "lib", "libgmpm_b200.so"), os.path.join(ROOT, "oracle", "libgmpm_oracle.so")]
    deps += [os.path.join(ROOT, "include", "multibody", "gpu_mpm", f) for f in os.listdir(os.path.join(ROOT, "include", "multibody", "gpu_mpm"))]
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= max(os.path.getmtime(d) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "oracle"), src, "-o", OUT, "-L" + os.path.join(ROOT, "drake_b200", "lib"),
           "-lgmpm_b200", "-L" + os.path.join(ROOT, "oracle"), "-lgmpm_oracle",
           "-Wl,-rpath,$ORIGIN/../../../drake_b200/lib:$ORIGIN/../../../oracle"]
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return OUT


OUT_TWO_RANK = os.path.join(HERE, "_build", "test_two_rank_host")


def build_two_rank(force=False):
    """tests/cpp/test_two_rank_host.cc: the multi-rank exchange driven from C++ through the peer-window ABI."""
    src = os.path.join(HERE, "test_two_rank_host.cc")
    deps = [src, os.path.join(ROOT, "drake_b200", "lib", "libgmpm_b200.so"), os.path.join(ROOT, "include", "gmpm_c.h")]
    if not force and os.path.exists(OUT_TWO_RANK) and os.path.getmtime(OUT_TWO_RANK) >= max(os.path.getmtime(d) for d in deps):
        return OUT_TWO_RANK
    os.makedirs(os.path.dirname(OUT_TWO_RANK), exist_ok=True)
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(cuda, "include"), src, "-o", OUT_TWO_RANK, "-L" + os.path.join(ROOT, "drake_b200", "lib"),
           "-lgmpm_b200", "-L" + os.path.join(cuda, "lib64"), "-lcudart",
           "-Wl,-rpath,$ORIGIN/../../../drake_b200/lib:" + os.path.join(cuda, "lib64")]
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return OUT_TWO_RANK


REF_DRIVER = "/root/reference/multibody/plant/deformable_driver.h"
EXCERPT = os.path.join(HERE, "_build", "driver_excerpt.inc")
OUT_EXCERPT = os.path.join(HERE, "_build", "test_driver_excerpt")


def build_excerpt(force=False):
    """The drop-in compile proof: deformable_driver.h:196-271 (the MPM method bodies) extracted from the reference
    tree WHERE IT LIES into the git-ignored _build/ directory and compiled against include/.  Only possible where
    /root/reference exists (the build container); the binary travels to the GPU box.  Returns the path or None."""
    src = os.path.join(HERE, "test_driver_excerpt.cc")
    if not os.path.exists(REF_DRIVER):
        return OUT_EXCERPT if os.path.exists(OUT_EXCERPT) else None
    deps = [src, REF_DRIVER, os.path.join(ROOT, "drake_b200", "lib", "libgmpm_b200.so")]
    deps += [os.path.join(ROOT, "include", "multibody", "gpu_mpm", f) for f in os.listdir(os.path.join(ROOT, "include", "multibody", "gpu_mpm"))]
    if not force and os.path.exists(OUT_EXCERPT) and os.path.getmtime(OUT_EXCERPT) >= max(os.path.getmtime(d) for d in deps):
        return OUT_EXCERPT
    os.makedirs(os.path.dirname(EXCERPT), exist_ok=True)
    lines = open(REF_DRIVER).read().split("\n")
    with open(EXCERPT, "w") as f:
        f.write("// generated at build time from %s lines 196-271 (not part of the repository)\n" % REF_DRIVER)
        f.write("\n".join(lines[195:271]) + "\n")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-Wno-unused-variable", "-Wno-unused-parameter", "-Wno-format",
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.dirname(EXCERPT), src, "-o", OUT_EXCERPT,
           "-L" + os.path.join(ROOT, "drake_b200", "lib"), "-lgmpm_b200", "-Wl,-rpath,$ORIGIN/../../../drake_b200/lib"]
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return OUT_EXCERPT


if __name__ == "__main__":
    print(build_excerpt(force=True))
    print(build_two_rank(force=True))
    print(build(force=True))
