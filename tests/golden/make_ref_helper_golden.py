#!/usr/bin/env python
"""Regenerates tests/golden/ref_helpers.json from the REFERENCE ITSELF.

Runs only in the build container, where /root/reference exists: `make -C oracle ref` compiles
oracle/ref_tu/ref_host_probe.cu, which #includes the reference's multibody/gpu_mpm/
cuda_mpm_kernels.cuh + math_tools.cuh where they lie and evaluates the reference's own
__host__ __device__ helpers (cell_index, morton_code, givens_QR, svd2x2, polar_decompose2x2,
fixed_corotated_PK1_2D, compute_dphi_dF, project_strain, inverse3, make_from_one_unit_vector) on the
host CPU for seeded inputs, at float and at double.  Values are stored as C99 hex floats (exact).
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    if not os.path.isdir("/root/reference/multibody/gpu_mpm"):
        sys.exit("needs /root/reference (build container only)")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/ref_host_probe"])
    out = subprocess.check_output([os.path.join(ROOT, "oracle", "_ref", "ref_host_probe")])
    with open(os.path.join(ROOT, "tests", "golden", "ref_helpers.json"), "wb") as f:
        f.write(out)
    print("wrote tests/golden/ref_helpers.json (%d bytes)" % len(out))


if __name__ == "__main__":
    main()
