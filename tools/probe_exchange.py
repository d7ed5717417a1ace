"""Per-kernel cost of the peer-window exchange with TWO states on ONE GPU (windows shared by raw pointer, so the
"peer" stores are local stores): separates what the exchange costs in instructions and fences from what it costs
on NVLink.  Development probe; `GMPM_HALO_UNFUSED=1` selects the separate send / receive kernels.

  python tools/probe_exchange.py [--sheets 17] [--steps 60]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from drake_b200 import scenes, slab  # noqa: E402
from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sheets", type=int, default=17)
    ap.add_argument("--steps", type=int, default=60)
    args = ap.parse_args()
    import torch
    dtype, dt = np.float64, 1e-3
    sheets = scenes.stacked_sheets(2 * args.sheets, dtype=dtype)
    parts = slab.partition_meshes(len(sheets), 2)
    sts = []
    for p in parts:
        st = GpuMpmState(dtype, domain_bits=8, dx=1.0 / 128)
        for i in p:
            st.AddQRCloth(*sheets[i])
        st.Finalize()
        sts.append(st)
    sv = GpuMpmSolver()

    def sync_all():
        for s in sts:
            sv.GpuSync(s)
        torch.cuda.synchronize()

    ex = slab.InProcessExchange([slab.GpuHaloBackend(s) for s in sts], sync=sync_all, peer=True)

    def step(i):
        for s in sts:
            sv.RebuildMapping(s, i % 40 == 0)
            sv.CalcFemStateAndForce(s, dt)
            sv.ParticleToGrid(s, dt)
        n = ex.exchange()
        for s in sts:
            sv.UpdateGrid(s, -1)
            sv.GridToParticle(s, dt)
        return n

    for i in range(10):
        n = step(i)
    sync_all()
    sts[0].set_profiling(True)
    for i in range(args.steps):
        n = step(10 + i)
    sync_all()
    rep = sts[0].profile_report()
    print("N per state = %d, shared blocks = %d, fused = %s" % (sts[0].n_particles(), n, "GMPM_HALO_UNFUSED" not in os.environ))
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        print("  %-24s %9.2f us/launch x %d" % (k, v[0] / v[1] * 1e3, v[1]))


if __name__ == "__main__":
    main()
