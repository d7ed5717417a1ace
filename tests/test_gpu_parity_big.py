"""GPU parity at the configurations the bench numbers are quoted on (VERDICT r01 "What's weak" 1, 4):

  * cfg-2, N = 1 006 468, the bench scene itself (34 stacked 100x100 sheets, cold): one substep and 20
    substeps against the CPU oracle (OpenMP port) AND against the reference's own kernels (oracle/_ref),
    fp64 at the north_star tolerances (1e-9 per step; 1e-7 after 20) and fp32 against the fp32 oracle and
    the precise fp32 reference build; integer results bit-exact.
  * cfg-4, N = 2 012 936 (68 sheets) with two box pads: the contact solve with max_it in {1, 3} against the
    oracle, and a frictionless full solve against the reference's own UpdateContact
    (cuda_mpm_solver.cu:215-621).

All through the C ABI.  Tolerances are norm-wise relative (max|a-b| / max|b|), as in test_gpu_parity.py.
"""
import numpy as np
import pytest

import oracle_py as O
import refgpu
from drake_b200 import scenes
from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs

from test_gpu_parity import assert_particles_close, rel

pytestmark = pytest.mark.gpu
DT = 1e-3
THREADS = __import__("os").cpu_count() or 1


def build(dtype, n_sheets, dz=0.02, with_ref=True):
    sheets = scenes.stacked_sheets(n_sheets, dz=dz, dtype=dtype)
    st, orc = GpuMpmState(dtype), O.OracleMpm(dtype)
    ref = refgpu.RefMpm(dtype) if with_ref and refgpu.available() else None
    for s in sheets:
        st.AddQRCloth(*s)
        orc.AddQRCloth(*s)
        if ref is not None:
            ref.AddQRCloth(*s)
    st.Finalize()
    orc.Finalize()
    if ref is not None:
        ref.Finalize()
    return st, GpuMpmSolver(), orc, ref


def step_all(st, sv, orc, ref, n, sort_first):
    for i in range(n):
        s = sort_first and i == 0
        sv.RebuildMapping(st, s)
        sv.CalcFemStateAndForce(st, DT)
        sv.ParticleToGrid(st, DT)
        sv.UpdateGrid(st, -1)
        sv.GridToParticle(st, DT)
        if s:
            orc.RebuildMapping(True)     # substep_omp recomputes the (unsorted) keys itself
        orc.substep_omp(DT, -1, THREADS)
        if ref is not None:
            ref.substep(DT, -1, s)
    if ref is not None:
        ref.GpuSync()


# fp32 bars: SURVEY 8c sets ~1e-5 per step against the non-fast-math build; positions and velocities are held to
# that.  20 steps: rounding differences grow with the stiffness of the sheets (E = 1e4-scale stress against a
# 1e-3-scale mass), measured 1e-4-level; bar 5e-4.
CASES = [(np.float64, 1e-9, 1e-7), (np.float32, 1e-5, 5e-4)]


@pytest.mark.parametrize("dtype,tol1,tol20", CASES)
def test_cfg2_1m_one_and_twenty_substeps_vs_oracle_and_reference_kernels(dtype, tol1, tol20):
    st, sv, orc, ref = build(dtype, 34)
    N = st.n_particles()
    assert N == 1006468 == orc.n_particles
    sizes = {"positions": 3 * N, "velocities": 3 * N, "affine_matrices": 9 * N, "volumes": N}
    refget = (lambda name: ref.read(name, sizes[name])) if ref is not None else None

    def check(tol, tag):
        # element-wise bar: 100 x the norm-wise one, every component against max(|itself|, 1e-4 x the field's scale)
        assert_particles_close(st.read, orc.read, tol, (tag, "ours vs CPU oracle"), elementwise_tol=100 * tol if dtype == np.float64 else np.inf)   # fp32: recorded, not asserted
        for name in ("pids", "index_mappings"):   # (sort_keys of the OpenMP port are one substep old: compared below)
            assert np.array_equal(st.read(name), orc.read(name)), (tag, name)
        if ref is not None:
            assert_particles_close(orc.read, refget, tol, (tag, "CPU oracle vs reference kernels"))
            assert_particles_close(st.read, refget, tol, (tag, "ours vs reference kernels"))
            assert np.array_equal(st.read("pids"), ref.read("pids", N, np.int32)), tag
            assert np.array_equal(st.read("index_mappings"), ref.read("index_mappings", N, np.int32)), tag

    step_all(st, sv, orc, ref, 1, sort_first=True)   # the bench loop sorts on its first substep
    check(tol1, "1 substep")
    # the sort itself, on the full scene: keys and permutation bit-equal to the reference's radix_sort
    if ref is not None:
        for name in ("sort_keys", "sort_ids"):
            assert np.array_equal(st.read(name), ref.read(name, N, np.uint32)), name
    step_all(st, sv, orc, ref, 19, sort_first=False)
    check(tol20, "20 substeps")
    # a re-sort after the scene has moved: integer results still bit-equal (cold scene: no cell-boundary ties)
    sv.RebuildMapping(st, True)
    orc.RebuildMapping(True)
    same = np.array_equal(np.sort(st.read("sort_keys")), np.sort(orc.read("sort_keys")))
    if dtype == np.float64:
        assert same, "fp64 positions within 1e-7 fell into different cells"
    if same:
        for name in ("sort_keys", "sort_ids", "pids", "index_mappings"):
            assert np.array_equal(st.read(name), orc.read(name)), name
    if ref is not None:
        ref.close()
    orc.close()
    st.Destroy()
    # the measured figures, for profiles/ (the GPU box returns gpurun_out/)
    import json
    import os
    import test_gpu_parity as tgp
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "parity_elementwise_%s.json" % np.dtype(dtype).name), "w") as f:
            json.dump([{"case": t, "field": n, "norm_wise": r, "element_wise": e, "floor": tgp.ELEMENTWISE_FLOOR}
                       for t, n, r, e in tgp.ELEMENTWISE_LOG], f, indent=1)
    tgp.ELEMENTWISE_LOG.clear()


def test_cfg2_1m_hundred_substeps_vs_oracle():
    """north_star's second fp64 bar AT THE HEADLINE SIZE (VERDICT r02 "What's weak" 1: it was held at 29 602 particles
    only): 1e-6 relative after 100 substeps of cfg-2, 1 006 468 particles, the bench's own loop (re-sort every 40th
    substep), against the CPU oracle; integer state bit-equal; the element-wise figure held to 100 x the bar."""
    dtype = np.float64
    st, sv, orc, _ = build(dtype, 34, with_ref=False)
    assert st.n_particles() == 1006468
    for i in range(100):
        s = i % 40 == 0
        sv.RebuildMapping(st, s); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT)
        sv.UpdateGrid(st, -1); sv.GridToParticle(st, DT)
        if s:
            orc.RebuildMapping(True)
        orc.substep_omp(DT, -1, THREADS)
    assert_particles_close(st.read, orc.read, 1e-6, "100 substeps at 1M", elementwise_tol=1e-4)
    for name in ("pids", "index_mappings"):
        assert np.array_equal(st.read(name), orc.read(name)), name
    import test_gpu_parity as tgp
    print("element-wise / norm-wise after 100 substeps:", [(n, r, e) for _, n, r, e in tgp.ELEMENTWISE_LOG])
    tgp.ELEMENTWISE_LOG.clear()
    orc.close()
    st.Destroy()


def pads(step=0):
    half = np.array([0.02, 0.10, 0.10])
    centers = [np.array([0.27, 0.5, 0.45]), np.array([0.73, 0.5, 0.45])]
    vel = [np.array([0.1, 0.0, 0.0]), np.array([-0.1, 0.0, 0.0])]
    out = []
    for b in range(2):
        c = centers[b] + vel[b] * DT * step
        out.append((c - half, c + half, b + 1, vel[b], c))
    return out


def pad_contacts(pos):
    parts = [scenes.box_contacts(pos, lo, hi, rigid_id=rid, rigid_v=v, rigid_p_WB=c) for lo, hi, rid, v, c in pads()]
    return tuple(np.concatenate([p[i] for p in parts]) for i in range(7))


@pytest.mark.parametrize("max_it,mu,exact", [(1, 1.0, False), (3, 1.0, False), (3, 0.3, True)])
def test_cfg4_2m_contact_iterations_vs_oracle(max_it, mu, exact):
    """cfg-4 shape: 68 sheets (N = 2 012 936), two moving pads, k = 1e6, d = 1e-5.  Two free substeps (strains
    develop), then the DeformableDriver substep with the first Newton iterations of the contact solve."""
    dtype = np.float64
    st, sv, orc, _ = build(dtype, 68, dz=0.01, with_ref=False)
    assert st.n_particles() == 2012936
    st.ReallocateExternelBodies(3)
    orc.ReallocateExternelBodies(3)
    step_all(st, sv, orc, None, 2, sort_first=True)
    sv.SyncParticleStateToCpu(st)
    pos = st.positions_host().copy()
    sv.RebuildMapping(st, False); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT); sv.UpdateGrid(st, -1)
    orc.RebuildMapping(False); orc.CalcFemStateAndForce(DT); orc.ParticleToGrid(DT); orc.UpdateGrid(-1)
    c = pad_contacts(pos)
    assert len(c[0]) > 10000
    pairs = MpmParticleContactPairs(dtype)
    pairs.extend(*c)
    sv.CopyContactPairs(st, pairs)
    orc.CopyContactPairs(*c)
    s_ours = sv.UpdateContact(st, 0, 0, DT, mu, 1e6, 1e-5, False, exact, max_it)
    s_orc = orc.UpdateContact(DT, mu, 1e6, 1e-5, exact, max_it)
    assert s_ours.iterations == s_orc.iterations == max_it
    assert s_ours.grid_dofs == s_orc.grid_dofs > 0
    assert s_ours.line_search_evals == s_orc.line_search_evals
    for name in ("contact_vel0", "contact_vel", "grid_momentum"):
        assert rel(st.read(name), orc.read(name)) <= 1e-9, name
    assert np.array_equal(st.read("contact_sort_keys"), orc.read("contact_sort_keys"))
    st.ExternelBodyForceToHost()
    tau, f = orc.ExternelBodyForceToHost()
    assert rel(st.external_forces_host().F_Bq_W_f, f.reshape(-1, 3), 1e-12) <= 1e-8
    assert rel(st.external_forces_host().F_Bq_W_tau, tau.reshape(-1, 3), 1e-12) <= 1e-8
    sv.GridToParticle(st, DT)
    orc.GridToParticle(DT)
    assert_particles_close(st.read, orc.read, 1e-9, "after the coupled substep")
    orc.close()
    st.Destroy()


@pytest.mark.skipif(not refgpu.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_cfg4_2m_frictionless_solve_vs_reference_update_contact():
    """The same scene against the reference's own UpdateContact (its iteration cap is a constant, so the
    comparison is a full solve; frictionless: the iteration path holds no near-tie decisions)."""
    dtype = np.float64
    sheets = scenes.stacked_sheets(68, dz=0.01, dtype=dtype)
    st, ref = GpuMpmState(dtype), refgpu.RefMpm(dtype)
    for s in sheets:
        st.AddQRCloth(*s)
        ref.AddQRCloth(*s)
    st.Finalize()
    ref.Finalize()
    st.ReallocateExternelBodies(3)
    ref.ReallocateExternelBodies(3)
    sv = GpuMpmSolver()
    N = st.n_particles()
    for i in range(2):
        sv.RebuildMapping(st, i == 0); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT)
        sv.UpdateGrid(st, -1); sv.GridToParticle(st, DT)
        ref.substep(DT, -1, i == 0)
    sv.SyncParticleStateToCpu(st)
    pos = st.positions_host().copy()
    sv.RebuildMapping(st, False); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT); sv.UpdateGrid(st, -1)
    ref.RebuildMapping(False); ref.CalcFemStateAndForce(DT); ref.ParticleToGrid(DT); ref.UpdateGrid(-1)
    c = pad_contacts(pos)
    n = len(c[0])
    pairs = MpmParticleContactPairs(dtype)
    pairs.extend(*c)
    sv.CopyContactPairs(st, pairs)
    ref.CopyContactPairs(*c)
    stats = sv.UpdateContact(st, 0, 0, DT, 0.0, 1e6, 1e-5, False, False)
    ref.UpdateContact(DT, 0.0, 1e6, 1e-5, False)
    assert stats.residual <= 1e-4 and stats.iterations >= 1
    assert rel(st.read("contact_vel0"), ref.read("contact_vel0", 3 * n)) <= 1e-9
    assert rel(st.read("contact_vel"), ref.read("contact_vel", 3 * n)) <= 1e-8
    st.ExternelBodyForceToHost()
    tau_r, f_r = ref.ExternelBodyForceToHost()
    assert rel(st.external_forces_host().F_Bq_W_f, f_r, 1e-12) <= 1e-7
    sv.GridToParticle(st, DT)
    ref.GridToParticle(DT)
    ref.GpuSync()
    for name, k in (("positions", 3), ("velocities", 3)):
        assert rel(st.read(name), ref.read(name, k * N)) <= 1e-8, name
    ref.close()
    st.Destroy()
