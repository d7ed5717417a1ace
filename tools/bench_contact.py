"""cfg-4 of BASELINE.json: 2M-particle cloth stack squeezed by two moving box pads, one
DeformableDriver substep per iteration through the C ABI: position snapshot -> 4 solver calls ->
contact pairs built on the HOST from positions_host() (as CalcMpmContactPairs does,
deformable_driver.h:120-194; analytic boxes stand in for the FCL query) -> CopyContactPairs ->
UpdateContact (the whole Jacobi-Newton solve in one cooperative kernel) -> GridToParticle.

Prints one JSON line: particle-steps/s of the coupled substep, Newton iterations, device time of the
solve.  Development / evidence tool (profiles/); bench.py is the driver contract.

  python tools/bench_contact.py [--sheets 68] [--steps 30] [--exact] [--f32]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from drake_b200 import scenes  # noqa: E402
from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs  # noqa: E402

DT = 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sheets", type=int, default=68)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--exact", action="store_true")
    ap.add_argument("--f32", action="store_true")
    ap.add_argument("--device-pairs", action="store_true",
                    help="generate the contact pairs on the device (gmpm_generate_contact_pairs) instead of on the host")
    ap.add_argument("--library-loop", action="store_true",
                    help="the whole coupled loop in one ABI call (gmpm_run_driver_substeps), pairs generated on the device")
    ap.add_argument("--query-every", type=int, default=1, help="contact_query_frequency of the library loop")
    ap.add_argument("--mu", type=float, default=1.0)
    ap.add_argument("--stiffness", type=float, default=1e6)
    ap.add_argument("--damping", type=float, default=1e-5)
    args = ap.parse_args()
    dtype = np.float32 if args.f32 else np.float64
    st = GpuMpmState(dtype)
    # 68 sheets 0.01 apart: the same footprint as cfg-2 at twice the particle count
    for pos, vel, tri in scenes.stacked_sheets(args.sheets, dz=0.68 / args.sheets, dtype=dtype):
        st.AddQRCloth(pos, vel, tri)
    st.Finalize()
    st.ReallocateExternelBodies(3)
    sv = GpuMpmSolver()
    N = st.n_particles()
    # two gripper pads (mpm_cloth_folding.cc:32-34 scaled to the stack): boxes closing in along x at 0.1 m/s
    half = np.array([0.02, 0.10, 0.10])
    centers = [np.array([0.27, 0.5, 0.45]), np.array([0.73, 0.5, 0.45])]
    pad_v = [np.array([0.1, 0.0, 0.0]), np.array([-0.1, 0.0, 0.0])]
    sv.RunSubsteps(st, 20, DT, -1, 40)   # settle
    sv.GpuSync(st)

    if args.library_loop:
        geoms = [{"kind": "box", "rigid_id": b + 1, "lo": centers[b] - half, "hi": centers[b] + half,
                  "velocity": pad_v[b], "p_WB": centers[b]} for b in range(2)]
        t_begin = time.perf_counter()
        ds = sv.RunDriverSubsteps(st, args.steps, DT, geoms, -1, 0, args.query_every, args.mu, args.stiffness,
                                  args.damping, args.exact)
        sv.GpuSync(st)
        wall = time.perf_counter() - t_begin
        st.ExternelBodyForceToHost()
        f = st.external_forces_host().F_Bq_W_f
        print(json.dumps({"workload": "cfg-4: %d sheets, N=%d, %s, two moving box pads, library-side driver loop, "
                                      "contact_query_frequency=%d" % (args.sheets, N, np.dtype(dtype).name, args.query_every),
                          "steps": args.steps, "particle_steps_per_sec_e2e": N * args.steps / wall,
                          "ms_per_substep_e2e": wall / args.steps * 1e3, "contact_queries": ds.contact_queries,
                          "contact_pairs_mean": ds.contact_pairs / max(1, ds.substeps),
                          "newton_iterations_mean": ds.newton_iterations / max(1, ds.substeps),
                          "newton_iterations_max": ds.newton_iterations_max,
                          "pad_impulse_x": [float(f[1][0]), float(f[2][0])]}))
        st.Destroy()
        return

    iters, n_pairs, solve_ms, host_ms = [], [], [], []
    t_begin = time.perf_counter()
    for i in range(args.steps):
        sv.SyncParticleStateToCpu(st)
        sv.RebuildMapping(st, False)
        sv.CalcFemStateAndForce(st, DT)
        sv.ParticleToGrid(st, DT)
        sv.UpdateGrid(st, -1)
        h0 = time.perf_counter()
        if args.device_pairs:
            geoms = []
            for b in range(2):
                c = centers[b] + pad_v[b] * DT * i
                geoms.append({"kind": "box", "rigid_id": b + 1, "lo": c - half, "hi": c + half, "velocity": pad_v[b], "p_WB": c})
            n_now = sv.GenerateContactPairs(st, geoms)
            host_ms.append((time.perf_counter() - h0) * 1e3)
        else:
            ph = st.positions_host()
            h0 = time.perf_counter()
            pairs = MpmParticleContactPairs(dtype)
            for b in range(2):
                c = centers[b] + pad_v[b] * DT * i
                pairs.extend(*scenes.box_contacts(ph, c - half, c + half, rigid_id=b + 1, rigid_v=pad_v[b], rigid_p_WB=c))
            host_ms.append((time.perf_counter() - h0) * 1e3)
            sv.CopyContactPairs(st, pairs)
            n_now = pairs.size()
        st.timer_mark(0)
        stats = sv.UpdateContact(st, 0, i, DT, args.mu, args.stiffness, args.damping, False, args.exact)
        st.timer_mark(1)
        sv.GridToParticle(st, DT)
        solve_ms.append(st.timer_elapsed_ms(0, 1))
        iters.append(stats.iterations)
        n_pairs.append(n_now)
    sv.GpuSync(st)
    wall = time.perf_counter() - t_begin
    st.ExternelBodyForceToHost()
    f = st.external_forces_host().F_Bq_W_f
    line = {"workload": "cfg-4: %d sheets, N=%d, %s, two moving box pads, mu=%g k=%g d=%g, %s line search"
                        % (args.sheets, N, np.dtype(dtype).name, args.mu, args.stiffness, args.damping,
                           "exact" if args.exact else "backtracking"),
            "steps": args.steps, "particle_steps_per_sec_e2e": N * args.steps / wall,
            "ms_per_substep_e2e": wall / args.steps * 1e3,
            "contact_pairs_mean": float(np.mean(n_pairs)), "newton_iterations_mean": float(np.mean(iters)),
            "newton_iterations_max": int(np.max(iters)), "solve_ms_mean": float(np.mean(solve_ms)),
            "solve_us_per_newton_iteration": float(1e3 * np.sum(solve_ms) / max(1, np.sum(iters))),
            "contact_pairs_from": "device (gmpm_generate_contact_pairs)" if args.device_pairs else "host (numpy analytic SDF on positions_host())",
            "contact_query_ms_mean": float(np.mean(host_ms)),
            "pad_impulse_x": [float(f[1][0]), float(f[2][0])]}
    print(json.dumps(line))
    st.Destroy()


if __name__ == "__main__":
    main()
