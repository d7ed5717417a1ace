"""The drop-in C++ headers (include/multibody/gpu_mpm/*): they compile as plain C++17 against
the C ABI, and the reference's smoke test + DeformableDriver substep written against them matches the oracle."""
import os
import subprocess
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp"))
import build_cpp  # noqa: E402


def test_shim_headers_compile_and_link():
    exe = build_cpp.build()
    assert os.path.exists(exe)


@pytest.mark.gpu
def test_driver_loop_through_cpp_shim_matches_oracle():
    exe = build_cpp.build()
    r = subprocess.run([exe, "40", "30"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("OK"), r.stdout


def test_reference_driver_method_bodies_compile_against_the_drop_in_headers():
    """deformable_driver.h:196-271 (InitalizeExternalContactForces, FinalizeExternalContactForces,
    CalcAbstractStates), verbatim from the reference tree, compile unchanged against include/ (the excerpt is
    extracted at build time where /root/reference exists; the binary travels)."""
    exe = build_cpp.build_excerpt()
    if exe is None:
        pytest.skip("no reference tree here and no prebuilt excerpt test")
    assert os.path.exists(exe)


@pytest.mark.gpu
def test_reference_driver_method_bodies_run_through_the_c_abi():
    exe = build_cpp.build_excerpt()
    if exe is None:
        pytest.skip("no reference tree here and no prebuilt excerpt test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "OK driver excerpt" in r.stdout, r.stdout


def test_two_rank_cpp_host_compiles():
    assert os.path.exists(build_cpp.build_two_rank())


@pytest.mark.gpu
def test_two_rank_cpp_host_matches_the_single_state():
    """INTEGRATION.md section 5 as a C++ program: two states coupled through gmpm_halo_peer_* (windows attached by
    pointer, flag all-reduce on the host) reproduce the single-state run."""
    exe = build_cpp.build_two_rank()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("OK two-rank"), r.stdout
