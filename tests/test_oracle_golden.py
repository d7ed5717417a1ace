"""Pins the CPU oracle against outputs of the REFERENCE ITSELF (tests/golden/ref_helpers.json, made by
tests/golden/make_ref_helper_golden.py from the reference's own __host__ __device__ helpers), and
against the SURVEY 8a known answers.  Bar: bit-exact, integers and floating point alike -- the
oracle restates the same expression trees and is built with -ffp-contract=off."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_py as O

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(HERE, "golden", "ref_helpers.json")) as f:
        return json.load(f)


def _arr(case, name, dt):
    return np.array([float.fromhex(x) for x in case[name]], dtype=dt)


def test_survey_known_answers():
    L = O.lib()
    # SURVEY.md 8a1 (probed from cuda_mpm_kernels.cuh:296-363 on host)
    assert L.ogmpm_cell_index(1, 2, 3) == 27
    assert L.ogmpm_cell_index(5, 9, 127) == 300887
    assert L.ogmpm_cell_index(64, 64, 96) == 1867776
    assert L.ogmpm_morton_code(31, 31, 31) == 32767


def test_cell_index_vs_reference(golden):
    L = O.lib()
    assert L.ogmpm_morton_code(31, 31, 31) == golden["morton_31_31_31"]
    for c in golden["cell_index"]:
        x, y, z = c["xyz"]
        key = L.ogmpm_cell_index(x, y, z)
        assert key == c["key"]
        assert L.ogmpm_morton_code(x >> 2, y >> 2, z >> 2) == c["block_morton"]
        inv = (C.c_uint32 * 3)()
        L.ogmpm_inverse_cell_index(key, inv)
        assert list(inv) == c["inv"] == [x, y, z]


@pytest.mark.parametrize("suf,dt", [("f32", np.float32), ("f64", np.float64)])
def test_fp_helpers_bit_exact_vs_reference(golden, suf, dt):
    L = O.lib()
    prm = O.default_params()
    f = lambda n: getattr(L, "ogmpm_%s_%s" % (n, suf))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for case in golden[suf]:
        F = _arr(case, "F", dt)
        Q, R = np.zeros(9, dt), np.zeros(9, dt)
        f("givens_qr33")(p(F), p(Q), p(R))
        assert np.array_equal(Q, _arr(case, "Q", dt)), "givens_QR<3,3> Q"
        assert np.array_equal(R, _arr(case, "R", dt)), "givens_QR<3,3> R"
        A2 = np.array([F[0], F[1], F[3], F[4]], dt)
        U, S, V = np.zeros(4, dt), np.zeros(4, dt), np.zeros(4, dt)
        f("svd2x2")(p(A2), p(U), p(S), p(V))
        assert np.array_equal(U, _arr(case, "U2", dt))
        assert np.array_equal(S, _arr(case, "S2", dt))
        assert np.array_equal(V, _arr(case, "V2", dt))
        P = np.zeros(9, dt)
        f("compute_dphi_dF")(C.byref(prm), p(F), p(P))
        assert np.array_equal(P, _arr(case, "P", dt)), "compute_dphi_dF"
        Fp = F.copy()
        f("project_strain")(C.byref(prm), p(Fp))
        assert np.array_equal(Fp, _arr(case, "Fproj", dt)), "project_strain"
        u = _arr(case, "u", dt)
        J = np.zeros(9, dt)
        f("make_from_one_unit_vector")(p(u), 2, p(J))
        assert np.array_equal(J, _arr(case, "J", dt)), "make_from_one_unit_vector"


def test_survey_fp_known_answers():
    # F=[1.1,.05,0;-.02,.95,.1;.03,0,.9] -> P[0]=4.122853e4 (f32), 4.122847e4 (f64);
    # project_strain leaves F[8]=0.899474
    L = O.lib()
    prm = O.default_params()
    for suf, dt, p0 in (("f32", np.float32, 4.122853e4), ("f64", np.float64, 4.122847e4)):
        F = np.array([1.1, .05, 0, -.02, .95, .1, .03, 0, .9], dt)
        P = np.zeros(9, dt)
        getattr(L, "ogmpm_compute_dphi_dF_" + suf)(C.byref(prm), F.ctypes.data_as(C.c_void_p),
                                                    P.ctypes.data_as(C.c_void_p))
        assert abs(P[0] - p0) / p0 < 2e-7
        Fp = F.copy()
        getattr(L, "ogmpm_project_strain_" + suf)(C.byref(prm), Fp.ctypes.data_as(C.c_void_p))
        assert abs(Fp[8] - 0.899474) < 1e-6


def test_stable_low16_sort_matches_numpy():
    L = O.lib()
    rng = np.random.default_rng(20261104)
    for n in (0, 1, 2, 127, 128, 129, 5000):
        keys = rng.integers(0, 1 << 21, size=n, dtype=np.uint32)
        ids = np.arange(n, dtype=np.uint32)
        ko, io = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        L.ogmpm_stable_sort_low_bits(keys.ctypes.data_as(C.c_void_p), ids.ctypes.data_as(C.c_void_p), n, 16,
                                     ko.ctypes.data_as(C.c_void_p), io.ctypes.data_as(C.c_void_p))
        perm = np.argsort(keys & 0xFFFF, kind="stable").astype(np.uint32)
        assert np.array_equal(io, perm)
        assert np.array_equal(ko, keys[perm])


def test_contact_terms_vs_sap_hunt_crossley_formulas():
    """Secondary pin: the lagged Hunt-Crossley impulse/gradient formulas of
    multibody/contact_solvers/sap/sap_hunt_crossley_constraint.cc:67-267, restated in double with
    numpy: gamma_n = k dt (phi0 - dt vn)(1 - d vn), dgamma_n/dvn, gamma_t = -mu n0 t_hat,
    G_t = mu n0 / ||vt||_s P_perp.  Also checks that grad is the negative derivative of the cost."""
    L = O.lib()
    prm = O.default_params()
    rng = np.random.default_rng(7)
    for _ in range(200):
        phi0 = float(rng.uniform(1e-4, 5e-3))
        dt, k, d, mu = 1e-3, float(rng.choice([1e5, 1e6])), float(rng.choice([1e-5, 0.1, 10.0])), float(rng.uniform(0, 1))
        v0 = rng.normal(0, 0.3, 3)
        v = rng.normal(0, 0.3, 3)
        H, G = np.zeros(9), np.zeros(3)
        pp = lambda a: a.ctypes.data_as(C.c_void_p)
        L.ogmpm_contact_grad_hess_f64(C.byref(prm), phi0, dt, k, d, mu, pp(v0), pp(v), pp(H), pp(G))
        v_hat = min(phi0 / dt, 1.0 / d)
        if v0[2] > v_hat:
            assert not H.any() and not G.any()
            continue
        eps = prm.epsv
        n0 = max(k * dt * phi0 * (1 - d * v0[2]), 0.0)
        ts = np.sqrt(v[0] ** 2 + v[1] ** 2 + eps ** 2)
        that = v[:2] / ts
        np.testing.assert_allclose(G[:2], -mu * n0 * that, rtol=1e-13, atol=1e-300)
        np.testing.assert_allclose(G[2], k * dt * (phi0 - dt * v[2]) * (1 - d * v[2]), rtol=1e-13)
        Pperp = np.eye(2) - np.outer(that, that)
        np.testing.assert_allclose(H.reshape(3, 3)[:2, :2], -mu * n0 / ts * Pperp, rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(H[8], k * dt * (-dt - d * phi0 + 2 * d * dt * v[2]), rtol=1e-13)
        # gradient consistency where the cost is smooth (vn below v_hat): -dl/dv == grad
        if v[2] < v_hat - 1e-3:
            h = 1e-6
            num = np.zeros(3)
            for a in range(3):
                vp, vm = v.copy(), v.copy()
                vp[a] += h
                vm[a] -= h
                lp = L.ogmpm_contact_cost_f64(C.byref(prm), phi0, dt, k, d, mu, pp(v0), pp(vp))
                lm = L.ogmpm_contact_cost_f64(C.byref(prm), phi0, dt, k, d, mu, pp(v0), pp(vm))
                num[a] = (lp - lm) / (2 * h)
            np.testing.assert_allclose(-num, G, rtol=2e-5, atol=1e-7 * max(1.0, abs(G).max()))
