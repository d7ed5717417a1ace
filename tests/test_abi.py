"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads, exports every symbol
include/gmpm_c.h declares, and fails LOUDLY (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libpath():
    from drake_b200 import build
    return build.build()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gmpm_c.h")).read()
    macro = text[text.index("#define GMPM_DECLARE"):text.index("GMPM_DECLARE(f32, float)")]
    typed = set(re.findall(r"gmpm_(\w+?)_##SUF\(", macro))
    rest = text[:text.index("#define GMPM_DECLARE")] + text[text.index("GMPM_DECLARE(f64, double)"):]
    plain = set(re.findall(r"\b(gmpm_\w+)\(", rest))
    return plain, typed


def test_header_symbols_match_binding_tables():
    from drake_b200 import _lib
    plain, typed = _declared_symbols()
    assert plain == set(_lib.PLAIN_SYMBOLS)
    assert typed == set(_lib.TYPED_SYMBOLS)


def test_library_exports_every_declared_symbol(libpath):
    plain, typed = _declared_symbols()
    out = subprocess.check_output(["nm", "-D", "--defined-only", libpath]).decode()
    exported = set(l.split()[-1] for l in out.splitlines() if l.strip())
    want = set(plain) | {"gmpm_%s_%s" % (t, s) for t in typed for s in ("f32", "f64")}
    assert want <= exported, sorted(want - exported)


def test_library_is_sm100a_only(libpath):
    out = subprocess.check_output(["cuobjdump", "-lelf", libpath]).decode()
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_library_loads_and_integer_helpers_match_survey(libpath):
    from drake_b200 import _lib
    L = _lib.lib()
    assert b"sm_100a" in L.gmpm_version()
    # SURVEY 8a1 known answers (reference cuda_mpm_kernels.cuh:296-363 probed on host)
    assert L.gmpm_cell_index(1, 2, 3) == 27
    assert L.gmpm_cell_index(5, 9, 127) == 300887
    assert L.gmpm_cell_index(64, 64, 96) == 1867776
    assert L.gmpm_morton_code(31, 31, 31) == 32767
    xyz = (C.c_uint32 * 3)()
    L.gmpm_inverse_cell_index(300887, xyz)
    assert list(xyz) == [5, 9, 127]


def test_integer_helpers_match_oracle(libpath):
    import oracle_py as O
    from drake_b200 import _lib
    L, OL = _lib.lib(), O.lib()
    rng = np.random.default_rng(1)
    for x, y, z in rng.integers(0, 512, size=(2000, 3)):
        assert L.gmpm_cell_index(int(x), int(y), int(z)) == OL.ogmpm_cell_index(int(x), int(y), int(z))


def test_no_cpu_fallback_without_device(libpath):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from drake_b200 import _lib
    from drake_b200.gpu_mpm import GpuMpmState, radix_sort_pairs
    assert _lib.lib().gmpm_device_count() == 0
    with pytest.raises(_lib.GmpmError) as e:
        GpuMpmState(np.float64)
    assert e.value.status == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(_lib.GmpmError):
        radix_sort_pairs(np.arange(4, dtype=np.uint32), np.arange(4, dtype=np.uint32), 16)


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from drake_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.lib()


def test_contact_pairs_container_mirrors_reference():
    from drake_b200.gpu_mpm import MpmParticleContactPairs
    p = MpmParticleContactPairs(np.float32)
    p.push_back(3, 1, -0.01, [0, 0, 1], [0.5, 0.5, 0.1], [0, 0, 0], [0, 0, 0])
    p.extend([4, 5], [1, 2], [-0.02, -0.03], np.eye(3)[:2], np.ones((2, 3)), np.zeros((2, 3)), np.zeros((2, 3)))
    assert p.size() == 3
    ids, rid, dist, nrm, pos, rv, rp = p.arrays()
    assert ids.dtype == np.uint32 and list(ids) == [3, 4, 5] and list(rid) == [1, 1, 2]
    assert dist.dtype == np.float32 and nrm.shape == (3, 3) and pos.shape == (3, 3)
    p.clear()
    assert p.size() == 0 and p.arrays()[3].shape == (0, 3)
