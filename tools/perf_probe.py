"""Quick per-kernel timing of the no-contact substep (development aid; bench.py is the contract)."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from drake_b200 import scenes
from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState


def main():
    dtype = np.float64 if "--f32" not in sys.argv else np.float32
    n_sheets = 34
    for a in sys.argv:
        if a.startswith("--sheets="):
            n_sheets = int(a.split("=")[1])
    big = "--grid256" in sys.argv   # the multi-GPU configuration: 256^3 cells, dx = 1/128, far octant
    huge = "--grid512" in sys.argv  # cfg-3 / cfg-5 of BASELINE.json: 2^9 grid bits (4 m domain at dx = 1/128), tall stacks
    dz = 0.02
    for a in sys.argv:
        if a.startswith("--dz="):
            dz = float(a.split("=")[1])
    params = {"domain_bits": 9, "dx": 1.0 / 128} if huge else ({"domain_bits": 8, "dx": 1.0 / 128} if big else {})
    st = GpuMpmState(dtype, **params)
    for pos, vel, tri in scenes.stacked_sheets(n_sheets, dz=dz, dtype=dtype):
        if big:
            pos = (pos.astype(np.float64) + np.array([1.0, 1.0, 0.68])).astype(dtype)
        st.AddQRCloth(pos, vel, tri)
    st.Finalize()
    sv = GpuMpmSolver()
    N = st.n_particles()
    dt = 1e-3
    sv.RunSubsteps(st, 10, dt, -1, 40)
    sv.GpuSync(st)
    if "--quick" in sys.argv:
        sv.RunSubsteps(st, 12, dt, -1, 40)
        sv.GpuSync(st)
        print("quick ok", N)
        return
    for steps in (40, 200):
        t0 = time.perf_counter()
        sv.RunSubsteps(st, steps, dt, -1, 40)
        sv.GpuSync(st)
        t1 = time.perf_counter()
        print("N=%d %s steps=%d: %.3f ms/step, %.3f G particle-steps/s (wall)" %
              (N, np.dtype(dtype).name, steps, (t1 - t0) / steps * 1e3, N * steps / (t1 - t0) / 1e9))
    print("pblocks, touched:", st.block_counts(), "device MB:", st.device_bytes() / 1e6, st.dev_counters())
    st.set_profiling(True)
    sv.RunSubsteps(st, 40, dt, -1, 40)
    sv.GpuSync(st)
    rep = st.profile_report()
    tot = sum(v[0] for v in rep.values())
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        print("  %-24s %9.3f ms total %6d launches %9.2f us/launch %5.1f%%" % (k, v[0], v[1], v[0] / v[1] * 1e3, 100 * v[0] / tot))
    print("  profiled total %.3f ms over 40 steps = %.3f ms/step" % (tot, tot / 40))


if __name__ == "__main__":
    main()
