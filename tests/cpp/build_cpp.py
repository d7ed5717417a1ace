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


if __name__ == "__main__":
    print(build(force=True))
