"""The friction solve's order-to-order spread, measured on the CPU oracle (no GPU needed), and -- on the GPU box --
the product's friction solve held to that measured spread (a13)."""
import numpy as np
import pytest

import contact_spread as CS
import refgpu
from drake_b200 import scenes
from test_gpu_contact import BOX, sheet_over_box

CASES = [(1.0, 1e-5, False), (0.3, 1e-2, True)]


def _problem(dtype=np.float64, res=40):
    import oracle_py as O
    sheets = sheet_over_box(dtype, res=res)
    s = O.OracleMpm(dtype)
    for pos, vel, tri in sheets:
        s.AddQRCloth(pos, vel, tri)
    s.Finalize()
    pos = s.read("positions").reshape(-1, 3)
    s.close()
    return sheets, scenes.box_contacts(pos, *BOX, rigid_id=1, rigid_v=(0.02, -0.01, 0.1))


@pytest.mark.parametrize("mu,d,exact", CASES)
def test_oracle_friction_solve_spread_under_contact_reordering(mu, d, exact):
    """Same pairs, same problem, different summation order: the converged contact velocities differ by far more
    than rounding (the stop test bounds the residual by 1e-4 only) but stay inside a few percent."""
    sheets, c = _problem(res=32)
    base, sp = CS.measure_spread(sheets, c, mu, 1e6, d, exact, n_perm=3)
    print("spread mu=%g d=%g exact=%s: %s (base iterations %d)" % (mu, d, exact, sp, base[3].iterations))
    assert base[3].residual <= 1e-4
    assert sp["contact_vel"] <= 5e-2 and sp["force"] <= 1e-3
    assert 0.3 <= sp["it_lo"] and sp["it_hi"] <= 3.0


@pytest.mark.gpu
@pytest.mark.parametrize("mu,d,exact", CASES)
def test_friction_solve_sits_inside_the_measured_spread(mu, d, exact):
    """Ours against the oracle's identity-order solve: the deviation must not exceed the spread the reference
    algorithm shows against ITSELF under reordering (x3: the spread is a maximum over few samples; floor 1e-6)."""
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs
    dtype = np.float64
    sheets, c = _problem(dtype)
    base, sp = CS.measure_spread(sheets, c, mu, 1e6, d, exact, n_perm=5)
    st = GpuMpmState(dtype)
    for pos, vel, tri in sheets:
        st.AddQRCloth(pos, vel, tri)
    st.Finalize()
    st.ReallocateExternelBodies(3)
    sv = GpuMpmSolver()
    sv.RebuildMapping(st, False); sv.CalcFemStateAndForce(st, CS.DT); sv.ParticleToGrid(st, CS.DT); sv.UpdateGrid(st, -1)
    pairs = MpmParticleContactPairs(dtype)
    pairs.extend(*c)
    sv.CopyContactPairs(st, pairs)
    stats = sv.UpdateContact(st, 0, 0, CS.DT, mu, 1e6, d, False, exact)
    assert stats.residual <= 1e-4
    dev_v = CS.rel(st.read("contact_vel").reshape(-1, 3), base[0])
    st.ExternelBodyForceToHost()
    dev_f = CS.rel(st.external_forces_host().F_Bq_W_f, base[1], 1e-12)
    ratio = stats.iterations / base[3].iterations
    print("ours vs base: contact_vel %.3e (spread %.3e), force %.3e (spread %.3e), iterations x%.2f (spread %.2f..%.2f)"
          % (dev_v, sp["contact_vel"], dev_f, sp["force"], ratio, sp["it_lo"], sp["it_hi"]))
    assert dev_v <= max(3 * sp["contact_vel"], 1e-6)
    assert dev_f <= max(3 * sp["force"], 1e-6)
    assert 0.5 * sp["it_lo"] <= ratio <= 2.0 * sp["it_hi"]
    st.Destroy()


@pytest.mark.gpu
@pytest.mark.skipif(not refgpu.available(), reason="oracle/_ref not built")
def test_reference_kernels_show_the_same_spread_against_themselves():
    """The reference's own UpdateContact (oracle/_ref) under reordering of the pairs: its spread against itself is
    of the size the oracle's is, and the oracle's identity-order solve sits inside it."""
    dtype = np.float64
    mu, d, exact = CASES[0]
    sheets, c = _problem(dtype)
    make = lambda: refgpu.RefMpm(dtype)
    base_r, sp_r = CS.measure_spread(sheets, c, mu, 1e6, d, exact, n_perm=4, make=make)
    base_o, sp_o = CS.measure_spread(sheets, c, mu, 1e6, d, exact, n_perm=4)
    print("reference kernels' spread:", sp_r, " oracle's:", sp_o)
    assert CS.rel(base_o[0], base_r[0]) <= max(3 * max(sp_r["contact_vel"], sp_o["contact_vel"]), 1e-6)
    assert CS.rel(base_o[1], base_r[1], 1e-12) <= max(3 * max(sp_r["force"], sp_o["force"]), 1e-6)
