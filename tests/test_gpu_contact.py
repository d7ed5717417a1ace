"""GPU parity of the MPM-rigid contact solve, through the C ABI.

Checker = the CPU oracle's restatement of cuda_mpm_solver.cu:194-621 and kernels K11-K19, itself pinned
here against the reference's own kernels rebuilt for sm_100 (oracle/_ref).  The solve is an iterative
Newton method whose only run-to-run difference is the summation order of floating-point atomics; the
stop test (sqrt(sum|Dir|^2)/DoFs <= 1e-4) is the same expression on both sides.  Bars: iteration count
equal, fp64 results within 1e-9 relative after one solve (1e-7 after 10 coupled substeps)."""
import numpy as np
import pytest

import oracle_py as O
import refgpu
from drake_b200 import scenes
from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs

pytestmark = pytest.mark.gpu
DT = 1e-3


def rel(a, b, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), floor, 1e-300))


def sheet_over_box(dtype, res=40, seed=5):
    """A wavy sheet whose lower part dips into a box collider (mpm_cloth --testcase 0 geometry, scaled)."""
    rng = np.random.default_rng(seed)
    pos, vel, tri = scenes.cloth_sheet(res, 0.3, 0.35, 0.35, 0.5, dtype)
    pos = pos.astype(np.float64)
    pos[:, 2] += 0.012 * np.sin(11 * pos[:, 0]) * np.cos(9 * pos[:, 1])
    vel = rng.normal(0, 0.2, pos.shape)
    vel[:, 2] -= 0.5
    return [(pos.astype(dtype), vel.astype(dtype), tri)]


BOX = (np.array([0.3, 0.3, 0.40]), np.array([0.7, 0.7, 0.502]))


class Pair:
    """Ours and the oracle stepped in lock-step through the DeformableDriver substep."""

    def __init__(self, dtype, sheets, n_bodies=3):
        self.dtype = np.dtype(dtype)
        self.st, self.orc = GpuMpmState(dtype), O.OracleMpm(dtype)
        for pos, vel, tri in sheets:
            self.st.AddQRCloth(pos, vel, tri)
            self.orc.AddQRCloth(pos, vel, tri)
        self.st.Finalize()
        self.orc.Finalize()
        self.sv = GpuMpmSolver()
        self.st.ReallocateExternelBodies(n_bodies)
        self.orc.ReallocateExternelBodies(n_bodies)

    def pre_contact(self):
        self.sv.SyncParticleStateToCpu(self.st)
        pos = self.st.positions_host().copy()
        self.sv.RebuildMapping(self.st, False)
        self.sv.CalcFemStateAndForce(self.st, DT)
        self.sv.ParticleToGrid(self.st, DT)
        self.sv.UpdateGrid(self.st, -1)
        self.orc.RebuildMapping(False)
        self.orc.CalcFemStateAndForce(DT)
        self.orc.ParticleToGrid(DT)
        self.orc.UpdateGrid(-1)
        return pos

    def contact(self, c, mu, k, d, exact, max_it=0):
        pairs = MpmParticleContactPairs(self.dtype)
        pairs.extend(*c)
        self.sv.CopyContactPairs(self.st, pairs)
        self.orc.CopyContactPairs(*[np.asarray(x) for x in c])
        s_ours = self.sv.UpdateContact(self.st, 0, 0, DT, mu, k, d, False, exact, max_it)
        s_orc = self.orc.UpdateContact(DT, mu, k, d, exact, max_it)
        return s_ours, s_orc

    def g2p(self):
        self.sv.GridToParticle(self.st, DT)
        self.orc.GridToParticle(DT)


def check_solve(p, s_ours, s_orc, tol, tag):
    assert s_ours.iterations == s_orc.iterations, (tag, s_ours.iterations, s_orc.iterations)
    assert s_ours.grid_dofs == s_orc.grid_dofs, (tag, s_ours.grid_dofs, s_orc.grid_dofs)
    assert s_ours.line_search_evals == s_orc.line_search_evals, (tag, s_ours.line_search_evals, s_orc.line_search_evals)
    assert abs(s_ours.residual - s_orc.residual) <= max(tol * 100, 1e-12) * max(abs(s_orc.residual), 1e-30) + 1e-14, tag
    for name in ("contact_vel0", "contact_vel", "grid_momentum"):
        r = rel(p.st.read(name), p.orc.read(name))
        assert r <= tol, (tag, name, r)
    assert np.array_equal(p.st.read("contact_sort_keys"), p.orc.read("contact_sort_keys")), tag
    p.st.ExternelBodyForceToHost()
    ext = p.st.external_forces_host()
    tau, f = p.orc.ExternelBodyForceToHost()
    assert rel(ext.F_Bq_W_f, f.reshape(-1, 3), 1e-12) <= tol * 10, (tag, "F_f")
    assert rel(ext.F_Bq_W_tau, tau.reshape(-1, 3), 1e-12) <= tol * 10, (tag, "F_tau")


CASES = [(0.0, 0.0), (1.0, 1e-5), (0.3, 1e-2)]


@pytest.mark.parametrize("exact", [False, True])
@pytest.mark.parametrize("mu,d", CASES)
@pytest.mark.parametrize("max_it", [1, 3])
def test_contact_newton_iterations_fp64_vs_oracle(exact, mu, d, max_it):
    """The first Newton iterations are a deterministic function of the inputs (no near-tie decisions yet):
    every observable of the solve matches the oracle to rounding."""
    p = Pair(np.float64, sheet_over_box(np.float64))
    pos = p.pre_contact()
    c = scenes.box_contacts(pos, *BOX, rigid_id=1, rigid_v=(0.02, -0.01, 0.1))
    assert len(c[0]) > 50
    s_ours, s_orc = p.contact(c, mu, 1e6, d, exact, max_it)
    assert s_orc.iterations == max_it and s_orc.grid_dofs > 0
    check_solve(p, s_ours, s_orc, 1e-9, "exact=%s mu=%g d=%g it=%d" % (exact, mu, d, max_it))
    # observable scratch grids (reference layout) of the last iteration
    for name in ("grid_Hess", "grid_Grad", "grid_Dir"):
        assert rel(p.st.read(name), p.orc.read(name), 1e-12) <= 1e-9, name
    p.g2p()
    for name in ("positions", "velocities"):
        assert rel(p.st.read(name), p.orc.read(name)) <= 1e-9, name


@pytest.mark.parametrize("exact", [False, True])
def test_contact_full_solve_frictionless_fp64(exact):
    """mu = 0: the iteration path has no near-tie decisions -> same iteration count, results to rounding."""
    p = Pair(np.float64, sheet_over_box(np.float64))
    pos = p.pre_contact()
    c = scenes.box_contacts(pos, *BOX, rigid_id=1, rigid_v=(0.02, -0.01, 0.1))
    s_ours, s_orc = p.contact(c, 0.0, 1e6, 0.0, exact)
    assert s_orc.iterations >= 2
    check_solve(p, s_ours, s_orc, 1e-9, "frictionless exact=%s" % exact)
    p.g2p()
    for name in ("positions", "velocities"):
        assert rel(p.st.read(name), p.orc.read(name)) <= 1e-9, name


@pytest.mark.parametrize("exact", [False, True])
@pytest.mark.parametrize("mu,d", CASES[1:])
def test_contact_full_solve_friction_fp64(exact, mu, d):
    """With friction the reference's own iteration path depends on the order of its floating-point atomics
    (the backtracking test `E1 <= E0` and the Newton/bisection switches sit on rounding-level ties near
    convergence): permuting the contact order in the oracle moves the iteration count by +-30 % and the
    solution by ~1e-3 (measured).  The stop test only bounds sqrt(sum|Dir|^2)/DoFs by 1e-4.
    So the full solve is checked at that resolution: it stops by the same criterion, and the converged
    energy, contact velocities and body impulses agree within the spread of the reference algorithm."""
    p = Pair(np.float64, sheet_over_box(np.float64))
    pos = p.pre_contact()
    c = scenes.box_contacts(pos, *BOX, rigid_id=1, rigid_v=(0.02, -0.01, 0.1))
    s_ours, s_orc = p.contact(c, mu, 1e6, d, exact)
    assert s_ours.residual <= 1e-4 and s_orc.residual <= 1e-4
    assert 0.5 * s_orc.iterations <= s_ours.iterations <= 2 * s_orc.iterations, (s_ours.iterations, s_orc.iterations)
    assert s_ours.grid_dofs in (s_orc.grid_dofs - 2, s_orc.grid_dofs - 1, s_orc.grid_dofs, s_orc.grid_dofs + 1,
                                s_orc.grid_dofs + 2)
    assert abs(s_ours.energy - s_orc.energy) <= 2e-3 * abs(s_orc.energy), (s_ours.energy, s_orc.energy)
    assert rel(p.st.read("contact_vel0"), p.orc.read("contact_vel0")) <= 1e-9
    assert rel(p.st.read("contact_vel"), p.orc.read("contact_vel")) <= 2e-2
    p.st.ExternelBodyForceToHost()
    tau, f = p.orc.ExternelBodyForceToHost()
    assert rel(p.st.external_forces_host().F_Bq_W_f, f.reshape(-1, 3)) <= 1e-4
    assert rel(p.st.external_forces_host().F_Bq_W_tau, tau.reshape(-1, 3)) <= 1e-4


def test_contact_coupled_substeps_two_bodies_fp64():
    p = Pair(np.float64, sheet_over_box(np.float64, res=32))
    lo, hi = BOX
    pad = (np.array([0.45, 0.45, 0.495]), np.array([0.55, 0.55, 0.53]))   # a second, moving body from above
    for step in range(10):
        pos = p.pre_contact()
        c1 = scenes.box_contacts(pos, lo, hi, rigid_id=1)
        c2 = scenes.box_contacts(pos, pad[0], pad[1] , rigid_id=2, rigid_v=(0.0, 0.0, -0.1))
        c = tuple(np.concatenate([a, b]) for a, b in zip(c1, c2))
        s_ours, s_orc = p.contact(c, 0.0, 1e6, 1e-5, step % 2 == 1)
        assert s_ours.iterations == s_orc.iterations, step
        p.g2p()
    for name in ("positions", "velocities", "affine_matrices"):
        floor = 128.0 * float(np.max(np.abs(p.orc.read("velocities")))) if name == "affine_matrices" else 0.0
        assert rel(p.st.read(name), p.orc.read(name), floor) <= 1e-7, name
    p.st.ExternelBodyForceToHost()
    tau, f = p.orc.ExternelBodyForceToHost()
    assert rel(p.st.external_forces_host().F_Bq_W_f, f.reshape(-1, 3), 1e-12) <= 1e-6


def test_contact_fp32_vs_oracle():
    p = Pair(np.float32, sheet_over_box(np.float32))
    pos = p.pre_contact()
    c = scenes.box_contacts(pos, *BOX, rigid_id=1)
    s_ours, s_orc = p.contact(c, 0.5, 1e5, 1e-5, False, max_it=2)
    assert s_ours.iterations == s_orc.iterations == 2
    for name in ("contact_vel0", "contact_vel", "grid_momentum"):
        assert rel(p.st.read(name), p.orc.read(name)) <= 2e-3, name


def test_contact_edge_cases():
    p = Pair(np.float64, sheet_over_box(np.float64, res=16), n_bodies=2)
    pos = p.pre_contact()
    # no pairs: UpdateContact is a no-op (cuda_mpm_solver.cu:217)
    empty = tuple(a[:0] for a in scenes.box_contacts(pos, *BOX))
    s_ours, s_orc = p.contact(empty, 1.0, 1e6, 0.0, False)
    assert s_ours.iterations == 0 and p.st.num_contacts() == 0
    # pairs whose body id is outside the body table are solved but apply no impulse
    c = scenes.box_contacts(pos, *BOX, rigid_id=7)
    s_ours, s_orc = p.contact(c, 0.0, 1e6, 0.0, False)
    check_solve(p, s_ours, s_orc, 1e-9, "body id out of table")
    # shrinking and growing the pair set re-uses the buffers; one pair only
    c1 = tuple(a[:1] for a in c)
    p.g2p()
    pos = p.pre_contact()
    s_ours, s_orc = p.contact(c1, 0.0, 1e6, 0.0, True)
    assert s_ours.iterations == s_orc.iterations
    assert rel(p.st.read("contact_vel"), p.orc.read("contact_vel")) <= 1e-9
    # out-of-range particle index is rejected, not dereferenced
    bad = list(c1)
    bad[0] = np.array([p.st.n_particles()], np.uint32)
    pairs = MpmParticleContactPairs(np.float64)
    pairs.extend(*bad)
    with pytest.raises(Exception):
        p.sv.CopyContactPairs(p.st, pairs)


@pytest.mark.skipif(not refgpu.available(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("exact", [False, True])
@pytest.mark.parametrize("mu", [0.0, 1.0])
def test_contact_oracle_pinned_against_reference_kernels(exact, mu):
    """The oracle's contact restatement vs the reference's own UpdateContact (fp64 instantiation).
    Frictionless: to rounding.  With friction: within the spread of the reference's own atomics order."""
    dtype = np.float64
    sheets = sheet_over_box(dtype)
    orc, ref = O.OracleMpm(dtype), refgpu.RefMpm(dtype)
    for pos, vel, tri in sheets:
        orc.AddQRCloth(pos, vel, tri)
        ref.AddQRCloth(pos, vel, tri)
    orc.Finalize()
    ref.Finalize()
    orc.ReallocateExternelBodies(3)
    ref.ReallocateExternelBodies(3)
    for s in (orc, ref):
        s.RebuildMapping(False)
        s.CalcFemStateAndForce(DT)
        s.ParticleToGrid(DT)
        s.UpdateGrid(-1)
    pos = orc.read("positions").reshape(-1, 3)
    c = scenes.box_contacts(pos, *BOX, rigid_id=1, rigid_v=(0.02, -0.01, 0.1))
    orc.CopyContactPairs(*c)
    ref.CopyContactPairs(*c)
    orc.UpdateContact(DT, mu, 1e6, 1e-5, exact)
    ref.UpdateContact(DT, mu, 1e6, 1e-5, exact)
    n = len(c[0])
    tol_v, tol_f = (1e-9, 1e-8) if mu == 0.0 else (2e-2, 1e-4)
    assert rel(orc.read("contact_vel0"), ref.read("contact_vel0", 3 * n)) <= 1e-9
    assert rel(orc.read("contact_vel"), ref.read("contact_vel", 3 * n)) <= tol_v
    tau_o, f_o = orc.ExternelBodyForceToHost()
    tau_r, f_r = ref.ExternelBodyForceToHost()
    assert rel(f_o.reshape(-1, 3), f_r, 1e-12) <= tol_f
    assert rel(tau_o.reshape(-1, 3), tau_r, 1e-12) <= tol_f
    ref.close()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_device_generated_pairs_equal_the_host_path(dtype):
    """gmpm_generate_contact_pairs against the host loop it replaces: positions_host() ->
    analytic signed-distance query per geometry and particle -> CopyContactPairs.  Same pairs, same order,
    same values; the solves that follow agree."""
    geoms = [{"kind": "box", "rigid_id": 1, "lo": BOX[0], "hi": BOX[1], "velocity": (0.0, 0.0, 0.05)},
             {"kind": "sphere", "rigid_id": 2, "center": (0.45, 0.5, 0.52), "radius": 0.05, "velocity": (0.0, 0.0, -0.1)}]

    def host_pairs(ph):
        out = [scenes.box_contacts(ph, BOX[0], BOX[1], rigid_id=1, rigid_v=(0.0, 0.0, 0.05))]
        c, r = np.array(geoms[1]["center"]), geoms[1]["radius"]
        x = np.asarray(ph, dtype).reshape(-1, 3)          # the device evaluates in the state's scalar type
        rel = x - c.astype(dtype)
        ln = np.sqrt((rel * rel).sum(axis=1, dtype=dtype)).astype(dtype)
        idx = np.nonzero(ln < dtype(r))[0]
        n = len(idx)
        grad = rel[idx] / ln[idx, None]
        out.append((idx.astype(np.uint32), np.full(n, 2, np.uint32), ln[idx] - dtype(r), -grad, x[idx],
                    np.tile(np.array([0.0, 0.0, -0.1]), (n, 1)), np.tile(c, (n, 1))))
        return out

    runs = []
    for device_side in (False, True):
        st = GpuMpmState(dtype)
        for pos, vel, tri in sheet_over_box(dtype):
            st.AddQRCloth(pos, vel, tri)
        st.Finalize()
        st.ReallocateExternelBodies(3)
        sv = GpuMpmSolver()
        for it in range(3):
            sv.RebuildMapping(st, it == 1)     # the snapshot follows the re-sort: pair indices are visible indices
            sv.SyncParticleStateToCpu(st)      # (upstream's driver never sorts inside this loop)
            sv.CalcFemStateAndForce(st, DT)
            sv.ParticleToGrid(st, DT)
            sv.UpdateGrid(st, -1)
            if device_side:
                n = sv.GenerateContactPairs(st, geoms)
            else:
                pairs = MpmParticleContactPairs(dtype)
                for c in host_pairs(st.positions_host()):
                    pairs.extend(*c)
                sv.CopyContactPairs(st, pairs)
                n = pairs.size()
            assert n > 50
            arrays = {k: st.read(k) for k in ("contact_mpm_id", "contact_rigid_id", "contact_dist", "contact_normal",
                                              "contact_pos", "contact_rigid_v", "contact_rigid_p")}
            stats = sv.UpdateContact(st, 0, it, DT, 0.3, 1e6, 1e-5, False, False, 3)
            sv.GridToParticle(st, DT)
            runs.append((n, arrays, stats.iterations, st.read("positions")))
    half = len(runs) // 2
    for (n_h, a_h, it_h, p_h), (n_d, a_d, it_d, p_d) in zip(runs[:half], runs[half:]):
        assert n_h == n_d and it_h == it_d
        for k in a_h:
            if k in ("contact_mpm_id", "contact_rigid_id"):
                assert np.array_equal(a_h[k], a_d[k]), k       # same pairs in the same order
            else:
                # two separate runs (vertex forces are summed with atomics) and FMA contraction of |x - c|:
                # equal to rounding
                assert rel(a_d[k], a_h[k], 1e-3) <= (1e-12 if dtype == np.float64 else 1e-4), k   # f32: one ulp of x ~ 0.5 on |dist| ~ 1e-3
        assert rel(p_d, p_h) <= (1e-12 if dtype == np.float64 else 1e-5)


def test_library_side_driver_loop_equals_the_call_by_call_loop():
    """gmpm_run_driver_substeps: the DeformableDriver substep loop in ONE ABI call, contact
    pairs regenerated on the device while the geometries move, against the same loop issued call by call.
    With contact_query_frequency = 3 the pairs are reused between queries."""
    dtype = np.float64
    n_steps, mu, k, d = 6, 0.3, 1e6, 1e-5
    box_v = np.array([0.0, 0.0, 0.05])

    def fresh():
        st = GpuMpmState(dtype)
        for pos, vel, tri in sheet_over_box(dtype):
            st.AddQRCloth(pos, vel, tri)
        st.Finalize()
        st.ReallocateExternelBodies(2)
        return st

    def geom(i):
        s = box_v * DT * i
        return [{"kind": "box", "rigid_id": 1, "lo": np.array(BOX[0]) + s, "hi": np.array(BOX[1]) + s, "velocity": box_v,
                 "p_WB": 0.5 * (np.array(BOX[0]) + np.array(BOX[1])) + s}]

    sv = GpuMpmSolver()
    a = fresh()
    iters = pairs = 0
    for i in range(n_steps):
        sort = i % 4 == 0
        sv.RebuildMapping(a, sort)         # the snapshot follows the re-sort: pair indices are visible indices, and
        sv.SyncParticleStateToCpu(a)       # the solve reads the order after it (upstream never sorts in this loop)
        sv.CalcFemStateAndForce(a, DT)
        sv.ParticleToGrid(a, DT)
        sv.UpdateGrid(a, -1)
        pairs += sv.GenerateContactPairs(a, geom(i))
        iters += sv.UpdateContact(a, 0, i, DT, mu, k, d, False, False, 5).iterations
        sv.GridToParticle(a, DT)
    b = fresh()
    st = sv.RunDriverSubsteps(b, n_steps, DT, geom(0), -1, 4, 1, mu, k, d, False, 5)
    assert st.substeps == n_steps and st.contact_queries == n_steps
    assert st.contact_pairs == pairs and pairs > 50 * n_steps
    assert st.newton_iterations == iters and st.newton_iterations_max <= 5
    for name in ("positions", "velocities", "affine_matrices"):
        assert rel(b.read(name), a.read(name)) <= 1e-11, name
    a.ExternelBodyForceToHost()
    b.ExternelBodyForceToHost()
    fa, fb = a.external_forces_host().F_Bq_W_f, b.external_forces_host().F_Bq_W_f
    assert np.max(np.abs(np.asarray(fa) - np.asarray(fb))) <= 1e-9 * max(1.0, np.max(np.abs(np.asarray(fa))))
    # contact_query_frequency = 3: queries at substeps 0, 3 and at the re-sort of substep 4
    c = fresh()
    st3 = sv.RunDriverSubsteps(c, n_steps, DT, geom(0), -1, 4, 3, mu, k, d, False, 5)
    assert st3.substeps == n_steps and st3.contact_queries == 3
    assert st3.contact_pairs > 0 and np.all(np.isfinite(c.read("positions")))
    # stale pairs for at most two substeps (vertices in mesh order: the visible orders of two runs may differ)
    assert rel(c.DumpCpuState()[0], a.DumpCpuState()[0]) <= 5e-2
    # no geometry: the plain loop
    e, f = fresh(), fresh()
    sv.RunSubsteps(e, 4, DT, -1, 2)
    st0 = sv.RunDriverSubsteps(f, 4, DT, (), -1, 2)
    assert st0.contact_queries == 0 and st0.contact_pairs == 0
    assert rel(f.read("positions"), e.read("positions")) <= 1e-12
    for s in (a, b, c, e, f):
        s.Destroy()


def test_device_pairs_follow_a_resort_issued_after_the_snapshot():
    """SyncParticleStateToCpu -> RebuildMapping(sort=true) -> ... -> GenerateContactPairs.  The re-sort
    renumbers the visible order after the snapshot was taken; the device-side query must still pair the particles that
    are inside the geometry (same physical pairs as with the snapshot taken after the re-sort)."""
    dtype = np.float64
    geoms = [{"kind": "box", "rigid_id": 1, "lo": BOX[0], "hi": BOX[1], "velocity": (0.0, 0.0, 0.05)}]
    got = []
    for sync_first in (True, False):
        st = GpuMpmState(dtype)
        for pos, vel, tri in sheet_over_box(dtype):
            st.AddQRCloth(pos, vel, tri)
        st.Finalize()
        st.ReallocateExternelBodies(2)
        sv = GpuMpmSolver()
        sv.RunSubsteps(st, 3, DT, -1, 0)          # particles move: the 16-bit sort below really permutes
        if sync_first:
            sv.SyncParticleStateToCpu(st)
            sv.RebuildMapping(st, True)
        else:
            sv.RebuildMapping(st, True)
            sv.SyncParticleStateToCpu(st)
        sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT); sv.UpdateGrid(st, -1)
        n = sv.GenerateContactPairs(st, geoms)
        assert n > 50
        pids = st.read("pids")
        got.append((pids[st.read("contact_mpm_id")], st.read("contact_pos").reshape(-1, 3)))
        # every pair's stored position is the visible particle's position
        vis_pos = st.read("positions").reshape(-1, 3)
        st.Destroy()
    assert np.array_equal(got[0][0], got[1][0])
    # two separate runs: the vertex forces of the constitutive kernel are summed with atomics -> equal to rounding
    assert np.max(np.abs(got[0][1] - got[1][1])) <= 1e-12


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_culled_snapshot_gives_the_host_query_the_same_pairs(dtype):
    """gmpm_set_query_bounds: only the particles inside the declared bounds cross PCIe; positions_host() of the others
    reads 1e30.  The host query (box_contacts, standing in for ComputeSignedDistanceToPoint) must build the same pairs
    from the culled snapshot as from the full one, substep after substep (particles enter and leave the bounds)."""
    st = GpuMpmState(dtype)
    for pos, vel, tri in sheet_over_box(dtype):
        st.AddQRCloth(pos, vel, tri)
    st.Finalize()
    sv = GpuMpmSolver()
    N = st.n_particles()
    pad = 0.004
    st.SetContactQueryBounds([(BOX[0] - pad, BOX[1] + pad)])
    seen = set()
    for it in range(6):
        sv.RebuildMapping(st, it == 3)
        full = np.zeros((N, 3), dtype)
        sv.SyncParticleStateToCpu(st, full)                  # blocking form: always the full snapshot
        sv.SyncParticleStateToCpu(st)                        # async form: culled
        ph = st.positions_host()
        inside = np.all((full >= (BOX[0] - pad).astype(dtype)) & (full <= (BOX[1] + pad).astype(dtype)), axis=1)
        assert 0 < inside.sum() < N
        assert np.array_equal(ph[inside], full[inside])
        assert np.all(ph[~inside] == dtype(1e30))
        per = 4 + 3 * np.dtype(dtype).itemsize        # the fetch is speculative: at least the candidates, at most
        assert 4 + int(inside.sum()) * per <= st.snapshot_d2h_bytes() <= 4 + (int(inside.sum()) * 9 // 8 + 200) * per
        a = scenes.box_contacts(ph, *BOX, rigid_id=1)
        b = scenes.box_contacts(full, *BOX, rigid_id=1)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        seen.add(int(inside.sum()))
        sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT); sv.UpdateGrid(st, -1); sv.GridToParticle(st, DT)
    assert len(seen) > 1                                     # the candidate set changed over the run
    st.SetContactQueryBounds([])                             # no rigid geometry: nothing but the count crosses
    sv.SyncParticleStateToCpu(st)
    assert np.all(st.positions_host() == dtype(1e30)) and st.snapshot_d2h_bytes() == 4
    st.SetContactQueryBounds(None)                           # back to the full copy
    sv.SyncParticleStateToCpu(st)
    assert np.array_equal(st.positions_host().reshape(-1), st.read("positions"))
    assert st.snapshot_d2h_bytes() == 3 * N * np.dtype(dtype).itemsize
    st.Destroy()
