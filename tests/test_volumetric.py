"""Particle-only (volumetric) API and materials -- SURVEY 8f-4, an EXTENSION: upstream has one entry point
(AddQRCloth) and one law (the QR cloth).  Parity is UNPINNED by upstream; the checker is the CPU restatement of the
published algorithms in oracle/gmpm_oracle_vol.inc, itself checked here against closed forms (no GPU needed), and the
CUDA path is held to it at the same bars as the cloth path (1e-9 per substep in fp64)."""
import ctypes as C

import numpy as np
import pytest

import oracle_py as O
from drake_b200 import scenes

E, NU = 1e5, 0.3
MU, LAM = E / (2 * (1 + NU)), E * NU / ((1 + NU) * (1 - 2 * NU))


def update(model, F, Cm=None, dt=0.0, dtype=np.float64, phi=30.0, coh=0.0):
    suf = "f64" if dtype == np.float64 else "f32"
    F = np.ascontiguousarray(F, dtype).reshape(9).copy()
    Cm = np.zeros(9, dtype) if Cm is None else np.ascontiguousarray(Cm, dtype).reshape(9)
    tau = np.zeros(9, dtype)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    ct = C.c_double if dtype == np.float64 else C.c_float
    getattr(O.lib(), "ogmpm_volumetric_update_" + suf)(O.OracleMpm.MODELS[model], E, NU, phi, coh, p(Cm), ct(dt), p(F), p(tau))
    return F.reshape(3, 3), tau.reshape(3, 3)


def rot(axis, ang):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-13), (np.float32, 3e-5)])
def test_svd3_reconstructs_and_is_a_pair_of_rotations(dtype, tol):
    suf = "f64" if dtype == np.float64 else "f32"
    rng = np.random.default_rng(1)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for it in range(500):
        F = (np.eye(3) + rng.normal(0, 0.5, (3, 3))).astype(dtype)
        if it % 7 == 0:
            F[:, 2] = F[:, 0] * 0.3        # rank deficient
        if it % 11 == 0:
            F = -F                         # inverted
        U, V, s = np.zeros(9, dtype), np.zeros(9, dtype), np.zeros(3, dtype)
        getattr(O.lib(), "ogmpm_svd3_" + suf)(p(F), p(U), p(s), p(V))
        U, V = U.reshape(3, 3).astype(np.float64), V.reshape(3, 3).astype(np.float64)
        assert np.abs(U @ np.diag(s) @ V.T - F).max() < tol * max(1.0, np.abs(F).max())
        assert np.abs(U @ U.T - np.eye(3)).max() < tol and np.abs(V @ V.T - np.eye(3)).max() < tol
        assert np.linalg.det(U) > 0.99 and np.linalg.det(V) > 0.99
        assert s[0] >= s[1] >= abs(s[2]) - tol
        assert np.abs(np.sort(np.abs(s))[::-1] - np.linalg.svd(F.astype(np.float64), compute_uv=False)).max() < 10 * tol


@pytest.mark.parametrize("model", ["fixed_corotated", "stvk_hencky", "drucker_prager"])
def test_rigid_rotation_is_stress_free_and_objective(model):
    R = rot([1, 2, 3], 0.7)
    F, tau = update(model, R)
    assert np.abs(tau).max() < 1e-9 * MU and np.abs(F - R).max() < 1e-14
    # objectivity: tau(R F) = R tau(F) R^T
    F0 = np.diag([0.95, 0.9, 0.97]) @ rot([0, 1, 0], 0.2)
    _, t0 = update(model, F0)
    _, t1 = update(model, R @ F0)
    assert np.abs(t1 - R @ t0 @ R.T).max() < 1e-9 * MU


def test_closed_form_stresses():
    s = np.array([1.1, 0.97, 1.02])
    Q = rot([1, 1, 0], 0.4)
    F = Q @ np.diag(s) @ rot([0, 0, 1], -0.3).T
    J = s.prod()
    _, tau = update("fixed_corotated", F)
    want = Q @ np.diag(2 * MU * (s - 1) * s + LAM * (J - 1) * J) @ Q.T
    assert np.abs(tau - want).max() < 1e-10 * MU
    _, tau = update("stvk_hencky", F)
    e = np.log(s)
    want = Q @ np.diag(2 * MU * e + LAM * e.sum()) @ Q.T
    assert np.abs(tau - want).max() < 1e-10 * MU
    # F update: F <- (I + dt C) F
    Cm = np.array([[0.1, 0.2, 0.0], [-0.2, 0.0, 0.3], [0.0, 0.1, -0.1]])
    Fn, _ = update("stvk_hencky", F, Cm, 1e-2)
    assert np.abs(Fn - (np.eye(3) + 1e-2 * Cm) @ F).max() < 1e-15


def test_drucker_prager_return_mapping():
    phi = 35.0
    alpha = np.sqrt(2 / 3) * 2 * np.sin(np.radians(phi)) / (3 - np.sin(np.radians(phi)))
    k = (3 * LAM + 2 * MU) / (2 * MU) * alpha

    def yield_fn(F):
        e = np.log(np.linalg.svd(F, compute_uv=False))
        dev = e - e.mean()
        return np.linalg.norm(dev) + k * e.sum(), e

    # expansion: projected to the tip, stress free
    F, tau = update("drucker_prager", 1.05 * rot([1, 0, 0], 0.3), phi=phi)
    assert np.abs(np.linalg.svd(F, compute_uv=False) - 1).max() < 1e-14 and np.abs(tau).max() < 1e-9 * MU
    # hydrostatic compression: inside the cone, untouched, pressure = (3 lambda + 2 mu) log s
    F, tau = update("drucker_prager", 0.9 * np.eye(3), phi=phi)
    assert np.abs(F - 0.9 * np.eye(3)).max() < 1e-15
    assert np.abs(tau - np.eye(3) * (3 * LAM + 2 * MU) * np.log(0.9)).max() < 1e-9 * MU
    # compression with large shear: returned ONTO the cone along the deviatoric direction, trace preserved
    F0 = np.diag([0.7, 1.2, 1.0]) @ rot([0, 0, 1], 0.5)
    y0, e0 = yield_fn(F0)
    assert y0 > 0
    F, tau = update("drucker_prager", F0, phi=phi)
    y1, e1 = yield_fn(F)
    assert abs(y1) < 1e-12 and abs(e1.sum() - e0.sum()) < 1e-12
    # small shear under compression: elastic, untouched
    F0 = 0.8 * np.diag([1.01, 1.0, 0.99])
    assert yield_fn(F0)[0] < 0
    F, _ = update("drucker_prager", F0, phi=phi)
    assert np.abs(F - F0).max() < 1e-14


def block_scene(model, dtype=np.float64, n_cells=6, warm=True, **mat):
    dx = 1.0 / 128
    lo = np.array([0.40, 0.42, 0.30])
    pos, vel, vol = scenes.particle_block(lo, lo + n_cells * dx, dtype=dtype, seed=7)
    if warm:
        rng = np.random.default_rng(3)
        vel = (rng.normal(0, 0.3, pos.shape) + np.array([0.2, -0.1, -1.0])).astype(dtype)
    return dict(pos=pos, vel=vel, volume=vol, model=model, youngs=E, poisson=NU, **mat)


@pytest.mark.parametrize("model", ["fixed_corotated", "stvk_hencky", "drucker_prager"])
def test_oracle_block_free_fall_conserves_momentum(model):
    """In free fall the grid transfer conserves linear momentum exactly: sum m v changes by m g dt per substep
    (internal stresses cancel), for every material."""
    orc = O.OracleMpm(np.float64)
    sc = block_scene(model)
    orc.AddParticles(**sc)
    orc.Finalize()
    mass = orc.read("volumes") * orc.params.density
    p0 = (mass[:, None] * orc.read("velocities").reshape(-1, 3)).sum(0)
    for _ in range(10):
        orc.substep(1e-4)
    p1 = (mass[:, None] * orc.read("velocities").reshape(-1, 3)).sum(0)
    want = p0 + np.array([0, 0, orc.params.gravity * 10 * 1e-4 * mass.sum()])
    assert np.abs(p1 - want).max() < 1e-10 * np.abs(mass[:, None] * orc.read("velocities").reshape(-1, 3)).sum()
    F = orc.read("particle_F").reshape(-1, 3, 3)
    assert np.isfinite(F).all() and np.abs(F - np.eye(3)).max() > 1e-6     # the material really deformed


# ---------------------------------------------------------------------------------------------- GPU parity
def make_pair(dtype, bodies, cloth=(), **params):
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    st, orc = GpuMpmState(dtype, **params), O.OracleMpm(dtype, O.default_params(**params))
    for s in (st, orc):
        for c in cloth:
            s.AddQRCloth(*c)
        for b in bodies:
            s.AddParticles(**b)
        s.Finalize()
    return st, GpuMpmSolver(), orc


def rel(a, b, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), floor, 1e-300))


def run_both(st, sv, orc, n, dt, sort_every=0):
    for i in range(n):
        s = sort_every > 0 and i % sort_every == 0
        sv.RebuildMapping(st, s); sv.CalcFemStateAndForce(st, dt); sv.ParticleToGrid(st, dt)
        sv.UpdateGrid(st, -1); sv.GridToParticle(st, dt)
        orc.substep(dt, -1, s)


def check(st, orc, tol, tag):
    vmax = float(np.max(np.abs(orc.read("velocities"))))
    for name, floor in (("positions", 0), ("velocities", 0), ("affine_matrices", vmax * 128), ("particle_F", 0)):
        r = rel(st.read(name), orc.read(name), floor)
        assert r <= tol, (tag, name, r)
    assert np.array_equal(st.read("pids"), orc.read("pids")), tag


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["fixed_corotated", "stvk_hencky", "drucker_prager"])
def test_volumetric_block_vs_oracle_fp64(model):
    st, sv, orc = make_pair(np.float64, [block_scene(model, friction_angle_deg=30.0) if model == "drucker_prager"
                                         else block_scene(model)])
    assert st.n_volumetric() == st.n_particles() == orc.n_particles > 1500
    dt = 1e-4
    # stage by stage on a substep with real strains
    run_both(st, sv, orc, 3, dt)
    check(st, orc, 1e-9, "3 substeps")
    sv.RebuildMapping(st, False); orc.RebuildMapping(False)
    sv.CalcFemStateAndForce(st, dt); orc.CalcFemStateAndForce(dt)
    assert rel(st.read("taus"), orc.read("taus")) <= 1e-9
    assert rel(st.read("particle_F"), orc.read("particle_F")) <= 1e-12
    sv.ParticleToGrid(st, dt); orc.ParticleToGrid(dt)
    for name in ("grid_masses", "grid_momentum"):
        assert rel(st.read(name), orc.read(name)) <= 1e-9, name
    sv.UpdateGrid(st, -1); orc.UpdateGrid(-1)
    sv.GridToParticle(st, dt); orc.GridToParticle(dt)
    check(st, orc, 1e-9, "stage by stage")
    run_both(st, sv, orc, 60, dt, sort_every=16)       # falls, hits the floor region, deforms
    check(st, orc, 1e-6, "64 substeps")
    st.Destroy()


@pytest.mark.gpu
def test_cloth_and_two_volumetric_bodies_in_one_state_fp64():
    """A sheet falling onto a sand pile next to an elastic block: the three particle kinds share the grid."""
    dtype = np.float64
    sheet = scenes.cloth_sheet(30, 0.12, 0.36, 0.38, 0.40, dtype)
    sheet = (sheet[0], sheet[1] + np.array([0.0, 0.0, -0.8]), sheet[2])
    a = block_scene("drucker_prager", friction_angle_deg=35.0, cohesion=0.0)
    b = block_scene("fixed_corotated")
    b["pos"] = b["pos"] + np.array([0.06, 0.0, 0.0])
    st, sv, orc = make_pair(dtype, [a, b], cloth=[sheet])
    assert st.n_volumetric() == len(a["pos"]) + len(b["pos"]) and st.n_faces() > 0
    run_both(st, sv, orc, 1, 2e-4, sort_every=1)
    check(st, orc, 1e-9, "1 substep")
    run_both(st, sv, orc, 40, 2e-4, sort_every=9)
    check(st, orc, 1e-6, "41 substeps")
    # checkpoint round trip keeps the deformation gradients
    blob = st.save_state()
    st2, _, _ = make_pair(dtype, [a, b], cloth=[sheet])
    st2.load_state(blob)
    assert np.array_equal(st2.read("particle_F"), st.read("particle_F"))
    sv.RunSubsteps(st, 5, 2e-4, -1, 0); sv.RunSubsteps(st2, 5, 2e-4, -1, 0)
    assert rel(st2.read("positions"), st.read("positions")) <= 1e-11
    st.Destroy(); st2.Destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["fixed_corotated", "drucker_prager"])
def test_volumetric_fp32_vs_oracle(model):
    st, sv, orc = make_pair(np.float32, [block_scene(model, np.float32)])
    run_both(st, sv, orc, 20, 1e-4, sort_every=7)
    for name in ("positions", "velocities"):
        assert rel(st.read(name), orc.read(name)) <= 1e-3, name
    assert rel(st.read("particle_F"), orc.read("particle_F")) <= 1e-3
    st.Destroy()


@pytest.mark.gpu
def test_cfg2_named_workload_1m_fixed_corotated_block_properties():
    """BASELINE.json configs[1] as NAMED: a 1M-particle fixed-corotated elastic block drop, fp64, no rigid contact.
    Size-independent properties: mass on the grid = mass of the particles, momentum in free fall, det F > 0, and the
    first substeps against the oracle (OpenMP port)."""
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    dtype = np.float64
    dx = 1.0 / 128
    lo = np.array([0.3, 0.3, 0.35])
    pos, vel, vol = scenes.particle_block(lo, lo + 50 * dx, dtype=dtype)          # 100^3 particles
    assert len(pos) == 1000000
    st, orc = GpuMpmState(dtype), O.OracleMpm(dtype)
    for s in (st, orc):
        s.AddParticles(pos, vel, vol, "fixed_corotated", E, NU)
        s.Finalize()
    sv = GpuMpmSolver()
    dt = 1e-4
    import os
    for i in range(3):
        sv.RebuildMapping(st, i == 0); sv.CalcFemStateAndForce(st, dt); sv.ParticleToGrid(st, dt)
        if i == 2:
            mass = st.read("volumes") * st.params.density
            assert abs(st.read("grid_masses").sum() - mass.sum()) <= 1e-10 * mass.sum()
        sv.UpdateGrid(st, -1); sv.GridToParticle(st, dt)
        if i == 0:
            orc.RebuildMapping(True)
        orc.substep_omp(dt, -1, os.cpu_count() or 1)
    check(st, orc, 1e-9, "3 substeps at 1M")
    sv.RunSubsteps(st, 40, dt, -1, 20)
    F = st.read("particle_F").reshape(-1, 3, 3)
    assert np.isfinite(F).all() and np.linalg.det(F).min() > 0.5
    mom = (mass[:, None] * st.read("velocities").reshape(-1, 3)).sum(0)
    want_z = st.params.gravity * 43 * dt * mass.sum()
    assert abs(mom[2] - want_z) <= 1e-8 * abs(want_z) and np.abs(mom[:2]).max() <= 1e-8 * abs(want_z)
    st.Destroy()
