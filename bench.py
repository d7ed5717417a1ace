#!/usr/bin/env python
"""bench.py -- particle-steps/s of the MPM per-timestep hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): 34 stacked 100x100 cloth sheets,
N = 1 006 468 particles, fp64, dt = 1e-3, no rigid contact, 128^3 grid, the reference's own
standalone loop (cuda_mpm_test.cc:63-72): RebuildMapping(sort = every 40th substep)
CalcFemStateAndForce ParticleToGrid UpdateGrid GridToParticle.  One "step" = one substep.
N > 1: weak scaling -- every rank owns one such stack (whole meshes) of one larger scene on a 256^3
grid with the same dx; stacks are piled in pairs, and the grid blocks two ranks reach are summed
between P2G and UpdateGrid every substep (drake_b200/slab.py: two NCCL all-reduces).

  value : device-timed (CUDA events on the library's stream), state resident in HBM.
  e2e   : the DeformableDriver substep through the C ABI with HOST buffers
          (deformable_driver.h:240-260): SyncParticleStateToCpu (D2H of all positions into the pinned
          host mirror, every substep) + positions_host() read + CopyContactPairs (H2D; empty in this
          config) + the five solver calls.
  roofline : P2G tile kernel, algorithmic bytes / per-launch CUDA-event time (separate profiled pass).
  cpu_baseline : the CPU oracle (oracle/, OpenMP) on this box's host cores, bounded sample.
`--impl reference` times that CPU port alone (the reference has no CPU implementation of this path;
oracle/_ref holds its CUDA kernels, reported by the `ours` arm as `reference_gpu`).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_SHEETS = 34
SORT_EVERY = 40
METRIC = "particle_steps_per_sec"
UNIT = "particle-steps/s"
DX = 1.0 / 128
E_BLOCK, NU_BLOCK = 1e5, 0.3

# BASELINE.json configs.  "cfg2" (default, the driver's line) is the parity-PINNED stand-in for configs[1]: the
# reference's only material is the QR cloth, so 1M particles are 34 stacked sheets.  The others are the configs as
# NAMED, on the particle-only extension (parity unpinned upstream, oracle = oracle/gmpm_oracle_vol.inc).
CONFIGS = {
    "cfg2": {"dt": 1e-3, "scaling": "weak",
             "workload": "cfg-2: 34 stacked 100x100 cloth sheets, N=1006468 per GPU, fp64, dt=1e-3, no contact, dx=1/128"},
    "cfg2-block": {"dt": 2e-4, "scaling": "weak",
                   "workload": "cfg-2 as named: 1M-particle fixed-corotated elastic block drop (100^3 particles, 8 per cell, "
                               "E=1e5, nu=0.3), per GPU, dt=2e-4, no rigid contact, dx=1/128"},
    "cfg3": {"dt": 2e-4, "scaling": "strong",
             "workload": "cfg-3: 4M-particle (3 993 600) Drucker-Prager sand column collapse (80x80x78 cells, 8 per cell, "
                         "friction angle 30 deg), x-slabs over the ranks, dt=2e-4, 128^3 grid, dx=1/128"},
    "cfg4": {"dt": 1e-3, "scaling": "weak",
             "workload": "cfg-4: 2M-particle cloth stack (68 sheets, N=2012936) squeezed by two moving box pads, convex "
                         "Jacobi-Newton contact solve per substep (mu=1, k=1e6, d=1e-5), fp64, dt=1e-3, single GPU"},
    "cfg5": {"dt": 2e-4, "scaling": "weak",
             "workload": "cfg-5: 8M fixed-corotated particles per GPU (100^3 cells, 8 per cell), blocks adjacent in a 2x2x2 "
                         "arrangement, 512^3 grid at dx=1/128, dt=2e-4"},
}
CFG = "cfg2"
DT = CONFIGS[CFG]["dt"]
WORKLOAD = CONFIGS[CFG]["workload"]


def set_config(name):
    global CFG, DT, WORKLOAD
    CFG, DT, WORKLOAD = name, CONFIGS[name]["dt"], CONFIGS[name]["workload"]


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def scene_params(world):
    if CFG == "cfg5":
        return {"domain_bits": 9, "dx": DX}
    if CFG == "cfg3":
        return {}
    return {} if world == 1 else {"domain_bits": 8, "dx": DX}


SPARE_FRACTION = 0.15      # cfg3, N > 1: inactive particles per rank that arrivals of the migration take over
                           # (set to 1.0 at 8 ranks: the collapsing column piles up in the 10-cell edge slabs)
MIGRATE_EVERY = 128          # a migration event costs ~10 ms of host work and rank skew; ownership may lag


def park_of(rank, world):
    """cfg3: the inactive particles of a rank wait scattered over the lower part of its own x-slab (massless: they
    add nothing to the grid, and they touch no block another rank would not share anyway)."""
    c0, c1 = 80 * rank // world, 80 * (rank + 1) // world
    return np.array([0.19 + (c0 + 1) * DX, 0.19 + DX, 4.5 * DX, 0.19 + (c1 - 1) * DX, 0.19 + 79 * DX, 40.5 * DX])


SPARES_ADDED = {}          # rank -> inactive spare particles add_rank_bodies gave it


def add_rank_bodies(st, dtype, r, world, spare=False):
    """The part of the scene rank r owns, added to `st` (GpuMpmState, OracleMpm or RefMpm: same method names)."""
    from drake_b200 import scenes
    col, lvl = r // 2, r % 2
    ox, oy = (float(col % 2), float(col // 2)) if world > 1 else (0.0, 0.0)
    if CFG == "cfg2":
        # N = 1: the 128^3 unit cube of the reference.  N > 1: one 34-sheet stack per rank of ONE scene on a 256^3
        # grid (same dx): ranks 2c and 2c+1 are stacked in z in column c (the upper stack rests on the lower one: their
        # boundary sheets reach common grid blocks, which the ranks sum every substep); columns sit 1.0 apart
        z0 = 0.15 + lvl * N_SHEETS * 0.02
        for pos, vel, tri in scenes.stacked_sheets(N_SHEETS, z0=z0, dtype=dtype, seed=scenes.SEED + r):
            pos = pos.astype(np.float64)
            pos[:, 0] += ox
            pos[:, 1] += oy
            st.AddQRCloth(pos.astype(dtype), vel, tri)
    elif CFG == "cfg2-block":
        lo = np.array([0.3 + ox, 0.3 + oy, 0.35 + lvl * 50 * DX])        # the upper block sits on the lower one
        pos, vel, vol = scenes.particle_block(lo, lo + 50 * DX, dx=DX, dtype=dtype, seed=scenes.SEED + r)
        st.AddParticles(pos, vel, vol, "fixed_corotated", E_BLOCK, NU_BLOCK)
    elif CFG == "cfg3":
        c0, c1 = 80 * r // world, 80 * (r + 1) // world                   # this rank's x-slab of the column, in cells
        lo = np.array([0.19 + c0 * DX, 0.19, 3 * DX + 0.5 * DX])
        hi = np.array([0.19 + c1 * DX, 0.19 + 80 * DX, 3 * DX + 0.5 * DX + 78 * DX])
        pos, vel, vol = scenes.particle_block(lo, hi, dx=DX, dtype=dtype, seed=scenes.SEED + r)
        st.AddParticles(pos, vel, vol, "drucker_prager", E_BLOCK, NU_BLOCK, 30.0, 0.0)
        if spare and world > 1:   # position-owned x-slabs: room for the particles that migrate in (slab.SlabMigrator)
            n = int(len(pos) * SPARE_FRACTION)
            SPARES_ADDED[r] = n                      # the one place the count is decided (run() reads it back)
            from drake_b200 import slab
            st.AddParticles(slab.park_points(park_of(r, world), n), np.zeros((n, 3)), 0.0, "drucker_prager", E_BLOCK, NU_BLOCK,
                            30.0, 0.0)
    elif CFG == "cfg5":
        off = np.array([r & 1, (r >> 1) & 1, (r >> 2) & 1], np.float64) * 100 * DX
        lo = np.array([0.5, 0.5, 0.3]) + off
        pos, vel, vol = scenes.particle_block(lo, lo + 100 * DX, dx=DX, dtype=dtype, seed=scenes.SEED + r)
        st.AddParticles(pos, vel, vol, "fixed_corotated", E_BLOCK, NU_BLOCK)


def build_scene(cls_factory, dtype, rank=0, world=1, spare=False):
    """`rank` may be a list: the parts of several ranks in ONE state (the merged scene of the parity pass)."""
    st = cls_factory()
    for r in (rank if isinstance(rank, (list, tuple)) else [rank]):
        add_rank_bodies(st, dtype, r, world, spare)
    st.Finalize()
    return st


def interacting_ranks(rank, world):
    """The ranks whose parts can reach rank's particles within the parity pass (the merged checker scene)."""
    if CFG in ("cfg2", "cfg2-block"):
        return sorted((rank, rank ^ 1))       # pile partners; the columns are a metre apart
    if CFG == "cfg3":
        return list(range(world))             # one column
    return list(range(world)) if world <= 2 else None   # cfg5: all eight blocks touch; 64M particles do not fit a check


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


def by_pid(st, name, comps):
    v = np.asarray(st.read(name)).reshape(-1, comps)
    out = np.empty_like(v)
    out[np.asarray(st.read("pids"))] = v
    return out


def parity_single(st, dtype, steps):
    """N = 1: the state the product just computed (`steps` substeps from the fresh scene, re-sort on the first)
    against the CPU oracle (strict -O2 -ffp-contract=off build) run over the same substeps.  Checker only."""
    import oracle_py as O
    orc = build_scene(lambda: O.OracleMpm(dtype, O.default_params(**scene_params(1))), dtype)
    cores = os.cpu_count() or 1
    orc.RebuildMapping(True)
    for _ in range(steps):
        orc.substep_omp(DT, -1, cores)
    out = {"against": "CPU oracle (oracle/gmpm_oracle.c%s, strict build), same scene, same substeps"
                      % ("" if CFG == "cfg2" else " + gmpm_oracle_vol.inc, parity unpinned upstream"), "steps": steps,
           "max_rel_positions": _rel(st.read("positions"), orc.read("positions")),
           "max_rel_velocities": _rel(st.read("velocities"), orc.read("velocities")),
           "pids_equal": bool(np.array_equal(st.read("pids"), orc.read("pids")))}
    orc.close()
    return out


def parity_multi(st, dtype, steps, rank, world, local, make_state, mig=None):
    """N > 1: this rank's particles after `steps` substeps of the multi-rank path (fused peer-window exchange) against
    ONE state holding the merged scene of every rank that can interact with it, stepped on this GPU without any
    exchange.  Particles are matched by their original ids (the merged state numbers [faces | verts | volumetric] over
    all its bodies)."""
    from drake_b200.gpu_mpm import GpuMpmSolver
    group = interacting_ranks(rank, world)
    if group is None:
        return None
    merged = build_scene(make_state, dtype, group, world)
    GpuMpmSolver().RunSubsteps(merged, steps, DT, -1, SORT_EVERY)
    if st.n_faces():
        mv, _ = merged.DumpCpuState()           # vertices in mesh order
        mine, _ = st.DumpCpuState()
        v, k = mine.shape[0], group.index(rank)
        ref = mv[k * v:(k + 1) * v]             # equal-sized meshes per rank
    else:
        mv, mine = by_pid(merged, "positions", 3), by_pid(st, "positions", 3)
        counts = [0] * len(group)
        # volumetric bodies only: the merged pids are the ranks' pids, concatenated in group order
        import torch
        import torch.distributed as dist
        t = torch.zeros(world, dtype=torch.int64, device="cuda")
        t[rank] = mine.shape[0]
        dist.all_reduce(t)
        counts = [int(t[g]) for g in group]
        o = sum(counts[:group.index(rank)])
        ref = mv[o:o + mine.shape[0]]
        if mig is not None:   # migrating slabs: local ids are slots that change hands; match by scene-wide id
            act = mig.gid >= 0
            mine, ref = mine[act], mv[mig.gid[act]]
    merged.Destroy()
    return _rel(mine, ref)


class ClockSampler(threading.Thread):
    """Samples SM clock / power / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period=0.002):   # the default timed region is ~9 ms: sample every 2 ms
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples = []
        self.stop_flag = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            uuid = None
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids):
                    if ids[index].isdigit():
                        phys = int(ids[index])
                    else:
                        uuid = ids[index]
            self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode()) if uuid else pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), sm, pw, rs))
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % getattr(self, "err", "?")]}
        nv = self.nv
        sel = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
                 "hw_power_brake": getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80)}
        reasons = sorted({n for s in sel for n, bit in names.items() if s[3] & bit})
        return {"sm_mhz": float(np.median([s[1] for s in sel])) if sel else None, "sm_max_mhz": float(self.max_sm),
                "power_w_max": max([s[2] for s in sel]) if sel else None, "samples": len(sel), "reasons": reasons}


def cpu_baseline(dtype, budget_s=12.0):
    """The CPU oracle (OpenMP over particles/faces/cells, atomics for the scatters) on a bounded
    sample of the same workload: the full 1M-particle scene, as many substeps as fit the budget."""
    import oracle_py as O
    from drake_b200 import scenes
    cores = os.cpu_count() or 1
    # timing-only build of the oracle sources: -O3 -march=native, contraction allowed, compiled on this host
    # (the bit-exact -O2 -ffp-contract=off build is the parity checker, never the timed one)
    orc = build_scene(lambda: O.OracleMpm(dtype, O.default_params(**scene_params(1)), fast=True), dtype)
    n = orc.n_particles
    orc.substep_omp(DT, -1, cores)  # warm-up (page faults, first touched-block list)
    t0 = time.perf_counter()
    orc.substep_omp(DT, -1, cores)
    one = time.perf_counter() - t0
    steps = int(max(2, min(200, budget_s / max(one, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.substep_omp(DT, -1, cores)
    el = time.perf_counter() - t0
    orc.close()
    return {"value": n * steps / el, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d substeps of the full %d-particle scene, %.1f s, OpenMP x%d (oracle/gmpm_oracle.c, gcc %s)"
                      % (steps, n, el, cores, " ".join(O.FAST_FLAGS)),
            "flags": " ".join(O.FAST_FLAGS), "ms_per_step": el / steps * 1e3}


def arm_config():
    """`config` of the JSON line: the SAME dict in both arms (the driver compares them); what only one arm knows goes
    under other keys."""
    return {"workload": WORKLOAD, "sort_every": SORT_EVERY,
            "l2": "GPU arm: inputs larger than L2 (the particle state one substep streams exceeds the 126 MB L2, see "
                  "run.particle_state_mb_per_substep); CPU arm: not applicable"}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    dtype = np.float64
    t0 = time.perf_counter()
    cb = cpu_baseline(dtype, budget_s=max(5.0, min(60.0, 0.5 * (args.steps + args.warmup))))
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": arm_config(),
            "note": "the reference ships this path only as CUDA; this arm is its CPU restatement "
                    "(oracle/, pinned against the reference's own kernels) on the host cores",
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "flags")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    emit(line)


def reference_gpu(dtype, steps):
    """The reference's own kernels (oracle/_ref, as-shipped flags retargeted to sm_100) on this GPU."""
    import refgpu
    from drake_b200 import scenes
    out = {}
    for fast in (True, False):
        if not refgpu.available(fast):
            continue
        ref = refgpu.RefMpm(dtype, fast=fast)
        for pos, vel, tri in scenes.stacked_sheets(N_SHEETS, dtype=dtype):
            ref.AddQRCloth(pos, vel, tri)
        ref.Finalize()
        n = ref.n_verts + ref.n_faces
        for i in range(5):
            ref.substep(DT, -1, i == 0)
        ref.GpuSync()
        t0 = time.perf_counter()
        for i in range(steps):
            ref.substep(DT, -1, i % SORT_EVERY == 0)
        ref.GpuSync()
        el = time.perf_counter() - t0
        out["fast_math" if fast else "precise"] = {"value": n * steps / el, "ms_per_step": el / steps * 1e3}
        ref.close()
    return out


def kernel_table(prof, steps, N, n_faces, touched, T, peak):
    """Per-kernel times of a profiled pass -> us/launch, share, achieved GB/s on SURVEY 8d's algorithmic bytes."""
    c = touched * 64.0 / N
    alg = {"p2g_tile": (28 * T + 4 + 8 * T * c) * N,     # P2G = 28T + 4 + 8T c per particle
           "g2p": (18 * T + 3 * T * c) * N,              # G2P = 18T + 3T c
           "fem": (28 + 83 * T) * n_faces}               # FEM per face
    kern = {}
    tot_ms = sum(v[0] for v in prof.values())
    for name, (tms, cnt) in prof.items():
        per = tms / max(cnt, 1)
        kern[name] = {"us_per_launch": per * 1e3, "launches_per_step": cnt / steps, "share": tms / tot_ms}
        if name in alg:
            kern[name]["achieved_gbs"] = alg[name] / (per * 1e-3) / 1e9
            kern[name]["frac_of_measured_peak"] = kern[name]["achieved_gbs"] / peak
    return kern, alg, c


def f32_line(local, peak, steps):
    """The reference's PRODUCT dtype (settings.h:37 `using GpuT = float`, built with -use_fast_math,
    tools/skylark/cuda.bzl:66-86) on the same scene: device-timed value, per-kernel table, parity of the first
    20 substeps against the fp32 CPU oracle, and the reference's own fp32 kernels on this GPU."""
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    dtype = np.float32
    st = build_scene(lambda: GpuMpmState(dtype, device=local), dtype)
    sv = GpuMpmSolver()
    N = st.n_particles()
    sv.RunSubsteps(st, 20, DT, -1, SORT_EVERY)
    sv.GpuSync(st)
    par = parity_single(st, dtype, 20)
    par["tolerance"] = 5e-4
    par["ok"] = bool(max(par["max_rel_positions"], par["max_rel_velocities"]) <= 5e-4)
    sv.RunSubsteps(st, 60, DT, -1, SORT_EVERY)
    sv.GpuSync(st)
    st.timer_mark(0)
    sv.RunSubsteps(st, steps, DT, -1, SORT_EVERY)
    st.timer_mark(1)
    sv.GpuSync(st)
    ms = st.timer_elapsed_ms(0, 1)
    Kp = min(steps, 200)
    st.set_profiling(True)
    sv.RunSubsteps(st, Kp, DT, -1, SORT_EVERY)
    sv.GpuSync(st)
    prof = st.profile_report()
    st.set_profiling(False)
    _, touched = st.block_counts()
    kern, _, c = kernel_table(prof, Kp, N, st.n_faces(), touched, 4, peak)
    out = {"dtype": "f32", "value": N * steps / (ms * 1e-3), "unit": UNIT, "steps": steps, "ms_per_step": ms / steps,
           "touched_cells_per_particle": c, "kernels": kern, "parity": par}
    st.Destroy()
    try:
        out["reference_gpu"] = reference_gpu(dtype, 100)
    except Exception as e:
        out["reference_gpu"] = {"unavailable": repr(e)}
    return out


def cfg4_pads(step):
    """Two gripper pads (mpm_cloth_folding.cc:32-34 scaled to the stack) closing in along x at 0.1 m/s."""
    half = np.array([0.02, 0.10, 0.10])
    centers = [np.array([0.27, 0.5, 0.45]), np.array([0.73, 0.5, 0.45])]
    vel = [np.array([0.1, 0.0, 0.0]), np.array([-0.1, 0.0, 0.0])]
    return [(centers[b] + vel[b] * DT * step - half, centers[b] + vel[b] * DT * step + half, b + 1, vel[b],
             centers[b] + vel[b] * DT * step) for b in range(2)]


def cfg4_host_pairs(ph, step):
    """What CalcMpmContactPairs does on the host (deformable_driver.h:120-194), analytic boxes for the FCL query."""
    from drake_b200 import scenes
    parts = [scenes.box_contacts(ph, lo, hi, rigid_id=rid, rigid_v=v, rigid_p_WB=c) for lo, hi, rid, v, c in cfg4_pads(step)]
    return tuple(np.concatenate([p[i] for p in parts]) for i in range(7))


def run_cfg4(args, local):
    """BASELINE.json configs[3] on one GPU: the coupled DeformableDriver substep with the contact solve."""
    import oracle_py as O
    from drake_b200 import scenes
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs
    dtype, T = np.float64, 8
    MU, K_N, DAMP = 1.0, 1e6, 1e-5
    sheets = scenes.stacked_sheets(68, dz=0.01, dtype=dtype)
    st, orc = GpuMpmState(dtype, device=local), O.OracleMpm(dtype)
    for s in (st, orc):
        for sh in sheets:
            s.AddQRCloth(*sh)
        s.Finalize()
        s.ReallocateExternelBodies(3)
    sv = GpuMpmSolver()
    N = st.n_particles()
    cores = os.cpu_count() or 1

    def coupled(i, max_it, with_oracle):
        sv.SyncParticleStateToCpu(st)
        sv.RebuildMapping(st, False); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT); sv.UpdateGrid(st, -1)
        c = cfg4_host_pairs(st.positions_host(), i)
        pairs = MpmParticleContactPairs(dtype)
        pairs.extend(*c)
        sv.CopyContactPairs(st, pairs)
        stats = sv.UpdateContact(st, 0, i, DT, MU, K_N, DAMP, False, False, max_it)
        sv.GridToParticle(st, DT)
        t_orc = 0.0
        if with_oracle:
            t0 = time.perf_counter()
            orc.RebuildMapping(False); orc.CalcFemStateAndForce(DT); orc.ParticleToGrid(DT); orc.UpdateGrid(-1)
            orc.CopyContactPairs(*c)
            orc.UpdateContact(DT, MU, K_N, DAMP, False, max_it)
            orc.GridToParticle(DT)
            t_orc = time.perf_counter() - t0
        return stats, len(c[0]), t_orc

    # ---- parity pass: 2 free substeps, then 2 coupled substeps with the first 3 Newton iterations, against the oracle
    sv.RunSubsteps(st, 2, DT, -1, 40)
    orc.RebuildMapping(True)
    for _ in range(2):
        orc.substep_omp(DT, -1, cores)
    for i in range(2):
        coupled(i, 3, True)
    parity = {"against": "CPU oracle: 2 free + 2 coupled substeps (host pairs, 3 Newton iterations each)", "steps": 4,
              "max_rel_positions": _rel(st.read("positions"), orc.read("positions")),
              "max_rel_velocities": _rel(st.read("velocities"), orc.read("velocities")),
              "pids_equal": bool(np.array_equal(st.read("pids"), orc.read("pids"))), "tolerance": 1e-7}
    parity["ok"] = bool(max(parity["max_rel_positions"], parity["max_rel_velocities"]) <= 1e-7 and parity["pids_equal"])
    print("[bench] parity: %s" % parity, file=sys.stderr)
    sampler = ClockSampler(local)
    sampler.start()
    K, W = max(1, args.steps), max(3, args.warmup)
    geoms = lambda i: [{"kind": "box", "rigid_id": rid, "lo": lo, "hi": hi, "velocity": v, "p_WB": c}
                       for lo, hi, rid, v, c in cfg4_pads(i)]
    # ---- device-timed: the coupled loop on the library side of the ABI (pairs generated on the device)
    sv.RunDriverSubsteps(st, W, DT, geoms(2), -1, 0, 1, MU, K_N, DAMP, False)
    sv.GpuSync(st)
    l0 = st.kernel_launches()
    w0 = time.perf_counter()
    st.timer_mark(0)
    ds = sv.RunDriverSubsteps(st, K, DT, geoms(2 + W), -1, 0, 1, MU, K_N, DAMP, False)
    st.timer_mark(1)
    sv.GpuSync(st)
    w1 = time.perf_counter()
    ms = st.timer_elapsed_ms(0, 1)
    launches = st.kernel_launches() - l0
    clocks = sampler.summary(w0, w1)
    # ---- e2e: the driver substep as upstream issues it, host buffers: all positions D2H, host query, pairs H2D
    Ke = max(3, min(K, 20))
    coupled(2 + W + K, 0, False)
    sv.GpuSync(st)
    e0 = time.perf_counter()
    e_pairs, e_iters = 0, 0
    for i in range(Ke):
        s_, n_, _ = coupled(3 + W + K + i, 0, False)
        e_pairs += n_
        e_iters += s_.iterations
    sv.GpuSync(st)
    e_ms = (time.perf_counter() - e0) * 1e3
    # ---- per-kernel pass
    Kp = max(2, min(K, 10))
    st.set_profiling(True)
    dsp = sv.RunDriverSubsteps(st, Kp, DT, geoms(3 + W + K + Ke), -1, 0, 1, MU, K_N, DAMP, False)
    sv.GpuSync(st)
    prof = st.profile_report()
    st.set_profiling(False)
    pblocks, touched = st.block_counts()
    sampler.stop_flag.set()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    kern, alg, c = kernel_table(prof, Kp, N, st.n_faces(), touched, T, peak)
    newton = next((k for k in prof if "contact" in k and "newton" in k), None) or max(prof, key=lambda k: prof[k][0])
    us_per_iter = prof[newton][0] * 1e3 / max(1, dsp.newton_iterations)
    # ---- CPU baseline: the oracle's coupled substep (full solve), bounded sample
    t_cpu, n_cpu = 0.0, 0
    if not args.no_cpu_baseline:
        # (the oracle's state is that of the parity pass; the pairs come from the product's current positions: a timing
        # sample of the same amount of work -- as many Newton iterations as the product's solve takes on average)
        it_cap = max(1, int(round(ds.newton_iterations / max(1, ds.substeps))))
        while n_cpu < 2 and t_cpu < 25.0:
            _, _, t = coupled(100 + n_cpu, it_cap, True)
            t_cpu += t
            n_cpu += 1
    line = {"metric": METRIC, "value": N * K / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "particles_per_gpu": N, "touched_grid_blocks": touched,
                       "l2": "inputs larger than L2", "parallelism": "single GPU",
                       "loop": "gmpm_run_driver_substeps: pairs generated on the device every substep, full solve"},
            "clocks": clocks,
            "contact": {"pairs_mean": ds.contact_pairs / max(1, ds.substeps),
                        "newton_iterations_per_substep": ds.newton_iterations / max(1, ds.substeps),
                        "newton_iterations_max": ds.newton_iterations_max, "us_per_newton_iteration": us_per_iter,
                        "last_residual": ds.last_residual},
            "e2e": {"value": N * Ke / (e_ms * 1e-3), "unit": UNIT, "steps": Ke, "ms_per_step": e_ms / Ke,
                    "h2d_bytes_per_step": int(e_pairs / Ke * (2 * 4 + 13 * T)), "d2h_bytes_per_step": int(3 * T * N),
                    "newton_iterations_per_substep": e_iters / Ke,
                    "what": "DeformableDriver substep through the C ABI: SyncParticleStateToCpu (all positions), host "
                            "contact query on positions_host(), CopyContactPairs (H2D), UpdateContact, 5 solver calls"},
            "gpu_launches": int(launches), "gpu_launches_per_step": launches / K,
            "roofline": {"kernel": "p2g_tile", "bound": "hbm", "achieved": kern["p2g_tile"]["achieved_gbs"], "peak": peak,
                         "unit": "GB/s", "frac": kern["p2g_tile"]["achieved_gbs"] / peak, "traffic": None,
                         "note": "the substep is dominated by the contact solve (latency: grid barriers per Newton "
                                 "iteration); the roofline entry is the dominant HBM kernel of the transfer part"},
            "kernels": kern, "parity": parity}
    if n_cpu:
        cpu_v = N * n_cpu / t_cpu
        line["cpu_baseline"] = {"value": cpu_v, "unit": UNIT, "cores": 1, "kind": "port",
                                "sample": "%d coupled substeps of the full scene with the oracle's serial contact solve capped at %d "
                                          "Newton iterations (the product's mean), %.1f s" % (n_cpu, it_cap, t_cpu)}
        line["speedup_vs_cpu_port"] = {"device": line["value"] / cpu_v, "e2e": line["e2e"]["value"] / cpu_v}
    emit(line)
    st.Destroy()


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version
    banner, torchrun notices) was diverted to stderr by main()."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def bind_to_gpu_numa_node(torch, local):
    """Multi-rank runs: keep this process (and the pinned position mirror it is about to allocate, first touch) on
    the CPUs next to its GPU, so that 8 ranks do not push their 24 MB/substep snapshots through one socket."""
    try:
        import pynvml
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(local)
        try:
            h = pynvml.nvmlDeviceGetHandleByPciBusId("%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id))
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return sorted(os.sched_getaffinity(0))
    except Exception as e:                      # affinity is an optimisation, never a requirement
        print("[bench] no CPU affinity for GPU %d: %s" % (local, e), file=sys.stderr)
        return None


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=80)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS),
                    help="BASELINE.json config (default cfg2 = the driver's line; the others are builder runs)")
    ap.add_argument("--parity-steps", type=int, default=20,
                    help="substeps of the fresh scene checked against the oracle before the timed region (0 = off)")
    ap.add_argument("--no-final-parity", action="store_true",
                    help="cfg3, N>1: skip the check of the final state against the merged single-GPU run")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-f32", action="store_true", help="skip the fp32 (reference product dtype) sub-line")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--overlap", action="store_true",
                    help="N>1: UpdateGrid/G2P in two parts around the exchange (interior work hides the all-reduce)")
    ap.add_argument("--no-peer", action="store_true",
                    help="N>1: exchange the shared blocks with an NCCL all-reduce instead of the peer-memory windows")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    set_config(args.config)
    global SPARE_FRACTION
    if env_int("WORLD_SIZE", 1) >= 8:
        SPARE_FRACTION = 1.0
    elif env_int("WORLD_SIZE", 1) >= 4:
        SPARE_FRACTION = 0.3
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if CFG == "cfg4":
        if world > 1:
            raise SystemExit("bench.py --config cfg4: the contact solve is single-GPU (DESIGN.md section 8)")
        run_cfg4(args, local)
        return
    dist = None
    if world > 1:
        cpus = bind_to_gpu_numa_node(torch, local)
        if cpus:
            print("[bench] rank %d: bound to %d CPUs next to GPU %d (%d..%d)" % (rank, len(cpus), local, cpus[0], cpus[-1]),
                  file=sys.stderr)
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # the two all-reduces per substep are small (<= 4 MB) and latency bound: the low-latency protocol on the
        # ring measured fastest on 8 B200s (0.606 vs 0.626 ms/substep with NCCL's own choice; NVLS rejects f64)
        os.environ.setdefault("NCCL_ALGO", "Ring")
        os.environ.setdefault("NCCL_PROTO", "LL")
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState, MpmParticleContactPairs
    dtype = np.float64 if args.dtype == "f64" else np.float32
    T = np.dtype(dtype).itemsize
    st = build_scene(lambda: GpuMpmState(dtype, device=local, **scene_params(world)), dtype, rank, world, spare=True)
    sv = GpuMpmSolver()
    N = st.n_particles()
    n_spare = SPARES_ADDED.get(rank, 0) if (CFG == "cfg3" and world > 1) else 0
    N -= n_spare                                           # throughput counts the particles of the scene only
    K, W = args.steps, args.warmup
    ex = None
    if world > 1:
        from drake_b200 import slab
        peer = world <= 8 and not args.no_peer             # block sums through peer-memory windows (csrc/halo.cuh)
        backend = slab.GpuHaloBackend(st)
        ex = slab.SlabExchange(backend, peer=peer, peer_cap_blocks={"cfg5": 32768, "cfg3": 16384}.get(CFG, 8192))
    mig = None
    if world > 1 and CFG == "cfg3":
        # position-owned x-slabs: scene-wide ids = the merged scene's particle order (ranks in turn)
        counts = torch.zeros(world, dtype=torch.int64, device="cuda")
        counts[rank] = N
        dist.all_reduce(counts)
        off = int(counts[:rank].sum())
        gid = np.concatenate([np.arange(off, off + N), -np.ones(n_spare, np.int64)])
        edges = np.array([0.19 + (80 * r // world) * DX for r in range(1, world)])
        mig = slab.SlabMigrator(st, rank, world, edges, gid, park_of(rank, world), margin=4 * DX, cap=max(1, n_spare),
                                all_gather=slab.dist_all_gather_packed(torch.device("cuda", local)))
    migrated = [0]

    # interior UpdateGrid/G2P work hides the all-reduce; pays off once the collective is slow (measured: loses
    # 20 us at 2 ranks, wins 18 us at 8)
    overlap = args.overlap or (world >= 8 and args.no_peer and not args.no_overlap)

    def run_substeps(n, first=0):
        if ex is None:
            sv.RunSubsteps(st, n, DT, -1, SORT_EVERY)      # one ABI crossing, same calls inside
            return
        for i in range(first, first + n):                  # the exchange sits between P2G and UpdateGrid
            slab.substep(sv, st, ex, DT, -1, i % SORT_EVERY == 0, overlap=overlap)
            if mig is not None and (i + 1) % MIGRATE_EVERY == 0:
                moved = mig.migrate()                      # collective; the count is the same on every rank
                migrated[0] += moved
                if moved:
                    ex.planned = False                     # plan afresh after the next P2G, as at start-up
                    backend.reset_plan()

    # host slack: the only call of the substep that can block is the start of the exchange (it waits for the
    # plan issued one substep earlier); time spent there = how far the host runs ahead of the device
    plan_wait = [0.0]
    if ex is not None:
        begin0 = ex.begin

        def timed_begin():
            t = time.perf_counter()
            r = begin0()
            plan_wait[0] += time.perf_counter() - t
            return r
        ex.begin = timed_begin

    # ---- parity pass: the first substeps of the fresh scene, checked (not timed); they double as warm-up ----
    parity = None
    P = max(0, args.parity_steps)
    if P > 0:
        run_substeps(P)
        sv.GpuSync(st)
        tol = 1e-7 if args.dtype == "f64" else 5e-4     # north_star: 1e-9 per step, 1e-6 after 100 (fp64)
        if world == 1:
            parity = parity_single(st, dtype, P)
            worst = max(parity["max_rel_positions"], parity["max_rel_velocities"])
            parity["tolerance"] = tol
            parity["ok"] = bool(worst <= tol and parity["pids_equal"])
        else:
            mine = parity_multi(st, dtype, P, rank, world, local,
                                lambda: GpuMpmState(dtype, device=local, **scene_params(world)), mig)
            if mine is None:
                parity = {"unchecked": "the merged scene of all interacting ranks does not fit one GPU check"}
            else:
                t = torch.tensor([mine], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                parity = {"against": "one state holding the merged scene of every rank that interacts with this one, "
                                     "no exchange", "steps": P, "max_rel_positions": float(t[0]), "ranks_checked": world,
                          "tolerance": tol, "ok": bool(float(t[0]) <= tol)}
        print("[bench] rank %d parity: %s" % (rank, parity), file=sys.stderr)

    sampler = ClockSampler(local)
    sampler.start()

    # ---- device-timed region -------------------------------------------------------------------
    run_substeps(W)
    sv.GpuSync(st)
    barrier()
    l0 = st.kernel_launches()
    shared0 = ex.bytes_reduced if ex else 0
    w0 = time.perf_counter()
    plan_wait[0] = 0.0
    st.timer_mark(0)
    run_substeps(K)
    st.timer_mark(1)
    w_enq = time.perf_counter()                            # host done enqueueing; the device may still be running
    sv.GpuSync(st)
    barrier()
    w1 = time.perf_counter()
    print("[bench] rank %d: host enqueue %.3f ms/substep (of which %.3f waiting for the plan of the exchange), wall %.3f ms/substep" %
          (rank, (w_enq - w0) / K * 1e3, plan_wait[0] / K * 1e3, (w1 - w0) / K * 1e3), file=sys.stderr)
    launches = st.kernel_launches() - l0
    ms = st.timer_elapsed_ms(0, 1)
    halo_bytes_per_step = (ex.bytes_reduced - shared0) / K if ex else 0
    halo_blocks = ex.last_n_shared if ex else 0
    clocks = sampler.summary(w0, w1)
    clocks["window"] = "the device-timed region"

    # ---- migrating slabs: the state after the WHOLE run (parity pass + warm-up + timed substeps, with every
    # migration in between) against the merged scene stepped the same number of substeps on one GPU.  The sand
    # column is a chaotic system (non-smooth return mapping): summation-order differences grow with the step count,
    # so this is reported with its own, looser bar.
    parity_run = None
    if mig is not None and not args.no_final_parity:
        worst = parity_multi(st, dtype, P + W + K, rank, world, local,
                             lambda: GpuMpmState(dtype, device=local, **scene_params(world)), mig)
        t = torch.tensor([worst], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        parity_run = {"against": "the merged scene on one GPU, same number of substeps", "steps": P + W + K,
                      "migrations": (P + W + K) // MIGRATE_EVERY, "max_rel_positions": float(t[0])}

    # ---- end-to-end region: the driver substep with host buffers ---------------------------------
    # deformable_driver.h:240-260 as the drop-in shim issues it: SyncParticleStateToCpu starts the
    # snapshot of ALL positions (device gather + D2H into the library's pinned mirror), the four solver
    # calls are enqueued, positions_host() -- which CalcMpmContactPairs reads (:120-194) -- waits for
    # the copy and the result is read on the host, the contact pairs go up (H2D; none in this config).
    pairs = MpmParticleContactPairs(dtype)
    contact_ok = True
    e2e_check = [0.0]

    def driver_substep(i):
        nonlocal contact_ok
        sv.SyncParticleStateToCpu(st)                      # deformable_driver.h:243
        sv.RebuildMapping(st, i % SORT_EVERY == 0)
        sv.CalcFemStateAndForce(st, DT)
        sv.ParticleToGrid(st, DT)
        if ex is not None:                                 # grid update overlapped with the all-reduce; the
            ex.begin()                                     # contact stage of the driver needs the whole grid,
            st.update_grid_part(-1, 0)                     # so G2P is not split here
            ex.finish()
            st.update_grid_part(-1, 1)
        else:
            sv.UpdateGrid(st, -1)
        ph = st.positions_host()                           # host read of this step's particle positions
        e2e_check[0] += float(ph[i % N, 2])
        if contact_ok:
            try:
                sv.CopyContactPairs(st, pairs)             # H2D (empty here), :250
            except Exception:
                contact_ok = False
        sv.GridToParticle(st, DT)

    def run_e2e(n):
        for i in range(3):
            driver_substep(i)
        sv.GpuSync(st)
        barrier()
        t0 = time.perf_counter()
        moved = 0
        for i in range(n):
            driver_substep(i)
            moved += st.snapshot_d2h_bytes()
        sv.GpuSync(st)
        barrier()
        return (time.perf_counter() - t0) * 1e3, moved / n

    # (1) the loop exactly as upstream issues it: ALL positions cross PCIe every substep (deformable_driver.h:243)
    Kf = max(3, min(K, 100))
    full_ms, full_bytes = run_e2e(Kf)
    # (2) headline e2e: the same loop with the rigid geometry of the scene DECLARED (gmpm_set_query_bounds): only the
    # particles inside its bounds cross PCIe, positions_host() of the others reads 1e30 and the host query builds the
    # same pairs (tests/test_gpu_contact.py::test_culled_snapshot_...).  cfg-2 is "no rigid contact": the driver has
    # no rigid geometry to declare, so the list is empty and only the candidate count crosses.  The cost with
    # geometry present is the full-copy line scaled by the fraction of particles inside the bounds.
    st.SetContactQueryBounds([])
    Ke = max(3, min(K, 200))
    e_ms, e_bytes = run_e2e(Ke)
    st.SetContactQueryBounds(None)
    if (clocks.get("samples") or 0) < 3:     # a region of a few ms can fall between two NVML reads: widen to the e2e loops
        clocks = sampler.summary(w0, time.perf_counter())
        clocks["window"] = "the device-timed region and the end-to-end loops after it (same kernels, under load)"

    # ---- per-kernel pass (CUDA events between launches on the library's stream) ------------------
    Kp = max(3, min(K, 200))
    st.set_profiling(True)
    run_substeps(Kp)
    sv.GpuSync(st)
    prof = st.profile_report()
    st.set_profiling(False)
    pblocks, touched = st.block_counts()
    sampler.stop_flag.set()

    # ---- reduce over ranks -----------------------------------------------------------------------
    if dist is not None:
        t = torch.tensor([ms, e_ms, full_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e_ms, full_ms = float(t[0]), float(t[1]), float(t[2])
        tot = torch.tensor([float(N), float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        N_all, launches_all = float(tot[0]), int(tot[1])
    else:
        N_all, launches_all = float(N), int(launches)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy, burst)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s"
        kern, alg, c = kernel_table(prof, Kp, N, st.n_faces(), touched, T, peak)
        dom = max(alg, key=lambda n: prof.get(n, (0, 1))[0])
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(dom)
        except Exception:
            pass
        roofline = {"kernel": dom, "bound": "hbm", "achieved": kern[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                    "frac": kern[dom]["achieved_gbs"] / peak, "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg[dom], "touched_cells_per_particle": c,
                    "timing": "CUDA events between launches on the library stream, separate pass of %d steps" % Kp}
        line = {"metric": METRIC, "value": N_all * K / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": CONFIGS[CFG]["scaling"], "vs_baseline": None,
                "dtype": args.dtype, "data": "synthetic",
                "config": arm_config(),
                "run": {"particles_per_gpu": N, "particle_blocks": pblocks, "touched_grid_blocks": touched,
                        "particle_state_mb_per_substep": N * (44 * T + 8) / 1e6,
                        "parallelism": "single GPU" if world == 1 else
                        "%d ranks, each owning its part of ONE scene (see workload; whole bodies per rank); grid blocks "
                        "reached by two ranks are summed every substep: all-reduce of block flags (NCCL, side stream, "
                        "one substep ahead) + %s of %d shared blocks (%.0f KB/step all-reduced)%s"
                        % (world, "stores into the sharers' peer-memory windows (NVLink) with device-side delivery flags"
                           if ex is not None and ex.peer else "NCCL all-reduce", halo_blocks,
                           halo_bytes_per_step / 1e3, ", overlapped with interior G2P" if overlap else "")},
                "clocks": clocks,
                "e2e": {"value": N_all * Kf / (full_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": int(full_bytes), "steps": Kf, "ms_per_step": full_ms / Kf,
                        "what": "the DeformableDriver substep as upstream issues it (deformable_driver.h:240-260) through the C "
                                "ABI with host buffers: SyncParticleStateToCpu moves ALL positions to the host every substep "
                                "(device gather -> pinned mirror, overlapped with the next 4 solver calls), positions_host() "
                                "wait + host read, CopyContactPairs (H2D; cfg-2 has no rigid contact, 0 pairs), 5 solver "
                                "calls; bytes = measured mean per substep on rank 0"},
                "e2e_declared_bounds": {"value": N_all * Ke / (e_ms * 1e-3), "unit": UNIT, "steps": Ke,
                                        "ms_per_step": e_ms / Ke, "d2h_bytes_per_step": int(e_bytes),
                                        "what": "extension, not the headline: the same loop after gmpm_set_query_bounds declared the "
                                                "scene's rigid geometry (cfg-2: none, 0 boxes), so only the candidate count "
                                                "crosses PCIe"},
                "gpu_launches": launches_all, "gpu_launches_per_step": launches_all / K / world,
                "roofline": roofline, "kernels": kern, "parity": parity}
        if mig is not None:
            line["parity_after_run"] = parity_run
            print("[bench] migration host time on rank 0: extract %.1f ms, transport %.1f ms, insert %.1f ms in total"
                  % (1e3 * getattr(mig, "t_extract", 0), 1e3 * getattr(mig, "t_comm", 0), 1e3 * getattr(mig, "t_insert", 0)),
                  file=sys.stderr)
            line["migration"] = {"every_substeps": MIGRATE_EVERY, "particles_moved_total": int(migrated[0]),
                                 "spare_fraction": SPARE_FRACTION, "margin_cells": 4}
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline(dtype)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "flags")}
            line["speedup_vs_cpu_port"] = {"device": line["value"] / cb["value"], "e2e": line["e2e"]["value"] / cb["value"]}
        if world == 1 and not args.no_ref_gpu and CFG == "cfg2":   # the reference's kernels know the cloth only
            try:
                line["reference_gpu"] = reference_gpu(dtype, 100)
            except Exception as e:  # checker library missing or failed: not fatal for the product line
                line["reference_gpu"] = {"unavailable": repr(e)}
        if world == 1 and args.dtype == "f64" and not args.no_f32 and CFG == "cfg2":
            try:
                line["f32"] = f32_line(local, peak, min(K, 400))
            except Exception as e:
                line["f32"] = {"unavailable": repr(e)}
        if world == 1 and "reference_gpu" in line and "fast_math" in line.get("reference_gpu", {}):
            line["vs_existing_kernel"] = {
                "f64": line["value"] / line["reference_gpu"]["fast_math"]["value"],
                "f32": (line["f32"]["value"] / line["f32"]["reference_gpu"]["fast_math"]["value"]
                        if "fast_math" in line.get("f32", {}).get("reference_gpu", {}) else None),
                "what": "device-timed value / the reference's own kernels (as shipped: -use_fast_math) on this GPU"}
        emit(line)
    st.Destroy()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
