"""Order-to-order spread of the reference's friction solve (VERDICT r01 Missing 4 / a13).

UpdateContact (cuda_mpm_solver.cu:215-621) sums per-contact terms with floating-point atomics: the order of
the sums changes from run to run on the GPU.  With friction the Newton path sits on rounding-level ties
(backtracking `E1 <= E0`, the bisection switches), so a different summation order moves the iteration count
and, because the stop test only bounds sqrt(sum|Dir|^2)/DoFs by 1e-4, the converged solution.  This helper
MEASURES that spread on the CPU oracle (serial sums in contact order) by permuting the order of the contact
pairs -- the same pairs, the same problem, a different summation order -- so that the tolerance of the
friction-solve parity tests is derived, not asserted.  Test infrastructure only.
"""
import numpy as np

import oracle_py as O

DT = 1e-3


def rel(a, b, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), floor, 1e-300))


def solve_with_order(sheets, contacts, perm, mu, k, d, exact, n_bodies=3, dtype=np.float64, make=None):
    """One pre-contact substep + one full friction solve with the contact pairs in order `perm`.
    Returns per-pair contact velocities in the ORIGINAL pair order, body impulses, stats."""
    s = (make or (lambda: O.OracleMpm(dtype)))()
    for pos, vel, tri in sheets:
        s.AddQRCloth(pos, vel, tri)
    s.Finalize()
    s.ReallocateExternelBodies(n_bodies)
    s.RebuildMapping(False)
    s.CalcFemStateAndForce(DT)
    s.ParticleToGrid(DT)
    s.UpdateGrid(-1)
    c = tuple(np.asarray(a)[perm] for a in contacts)
    s.CopyContactPairs(*c)
    st = s.UpdateContact(DT, mu, k, d, exact)
    n = len(perm)
    if hasattr(s, "n_contacts"):
        cv = s.read("contact_vel").reshape(n, 3)
    else:
        cv = s.read("contact_vel", 3 * n).reshape(n, 3)
    out = np.empty_like(cv)
    out[perm] = cv
    tau, f = s.ExternelBodyForceToHost()
    s.close()
    return out, np.asarray(f).reshape(-1, 3), np.asarray(tau).reshape(-1, 3), st


def measure_spread(sheets, contacts, mu, k, d, exact, n_perm=5, seed=11, make=None):
    """Runs the identity order and n_perm random orders; returns the base solution and the largest deviation
    of any permuted run from it (contact velocity, body force, iteration-count ratio)."""
    n = len(contacts[0])
    rng = np.random.default_rng(seed)
    base = solve_with_order(sheets, contacts, np.arange(n), mu, k, d, exact, make=make)
    spread = {"contact_vel": 0.0, "force": 0.0, "torque": 0.0, "it_lo": 1.0, "it_hi": 1.0}
    for _ in range(n_perm):
        run = solve_with_order(sheets, contacts, rng.permutation(n), mu, k, d, exact, make=make)
        spread["contact_vel"] = max(spread["contact_vel"], rel(run[0], base[0]))
        spread["force"] = max(spread["force"], rel(run[1], base[1], 1e-12))
        spread["torque"] = max(spread["torque"], rel(run[2], base[2], 1e-12))
        if base[3] is not None and run[3] is not None and base[3].iterations > 0:
            r = run[3].iterations / base[3].iterations
            spread["it_lo"], spread["it_hi"] = min(spread["it_lo"], r), max(spread["it_hi"], r)
    return base, spread
