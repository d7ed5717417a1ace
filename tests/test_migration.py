"""Position-owned slabs with particle migration: a volumetric body split over two
(three) states by x-slabs, particles handed over as they cross the slab edges, must reproduce the single-state run.
CPU: the oracle backend (in process and over gloo, world size 2).  GPU: several states on one device."""
import multiprocessing as mp
import os
import socket

import numpy as np
import pytest

import oracle_py as O
from drake_b200 import scenes, slab
from oracle_halo import OracleHaloBackend

DT, DX = 2e-4, 1.0 / 128
MAT = dict(model="drucker_prager", youngs=1e5, poisson=0.3, friction_angle_deg=30.0, cohesion=0.0)


def moving_block(dtype, cells=(10, 6, 6)):
    """A sand block sliding in +x with a shear: within ~60 substeps a good part of it crosses the slab edges."""
    lo = np.array([0.40, 0.45, 0.30])
    pos, vel, vol = scenes.particle_block(lo, lo + np.array(cells) * DX, dtype=dtype, seed=11)
    vel = np.zeros_like(pos, dtype=np.float64)
    vel[:, 0] = 2.0 + 6.0 * (pos[:, 2] - lo[2])
    vel[:, 2] = -0.5
    return pos, vel.astype(dtype), vol


def park_of(rank):
    """Parking box of rank's inactive particles: a slab of cells off to the side of the scene."""
    return np.array([4.5 * DX, 4.5 * DX, (4.5 + 12 * rank) * DX, 24.5 * DX, 24.5 * DX, (12.5 + 12 * rank) * DX])


def make_part(factory, dtype, pos, vel, vol, mine, rank, spare):
    st = factory()
    st.AddParticles(pos[mine], vel[mine], vol, **MAT)
    st.AddParticles(slab.park_points(park_of(rank), spare), np.zeros((spare, 3)), 0.0, **MAT)
    st.Finalize()
    gid = np.concatenate([np.nonzero(mine)[0], -np.ones(spare, np.int64)])
    return st, gid


def gathered(parts, migrators, n_total, name="positions", comps=3):
    """Scene-wide array by global particle id from the active particles of all parts."""
    out = np.full((n_total, comps), np.nan)
    for st, m in zip(parts, migrators):
        v = np.asarray(st.read(name), np.float64).reshape(-1, comps)
        pids = np.asarray(st.read("pids"))
        by_pid = np.empty_like(v)
        by_pid[pids] = v
        act = m.gid >= 0
        out[m.gid[act]] = by_pid[act]
    assert not np.isnan(out).any()
    return out


def run_split(make_state, make_backend, step, dtype, world, steps, every, sync=None):
    pos, vel, vol = moving_block(dtype)
    n = len(pos)
    edges = slab.slab_edges(pos, world)
    own = slab.owner_of(pos[:, 0], edges)
    parts, migs = [], []
    for r in range(world):
        st, gid = make_part(make_state, dtype, pos, vel, vol, own == r, r, spare=n // 2)
        parts.append(st)
        migs.append(slab.SlabMigrator(st, r, world, edges, gid, park_of(r), margin=0.5 * DX, cap=n))
    ex = slab.InProcessExchange([make_backend(s) for s in parts], sync=sync)
    moved = 0
    for it in range(steps):
        step(parts, ex, it)
        if (it + 1) % every == 0:
            moved += slab.migrate_in_process(migs, ex)
    return parts, migs, moved, (pos, vel, vol)


def oracle_step(parts, ex, it):
    for s in parts:
        s.RebuildMapping(it % 7 == 0); s.CalcFemStateAndForce(DT); s.ParticleToGrid(DT)
    ex.exchange()
    for s in parts:
        s.UpdateGrid(-1); s.GridToParticle(DT)


@pytest.mark.parametrize("world", [2, 3])
def test_oracle_slabs_with_migration_match_single_state(world):
    dtype = np.float64
    parts, migs, moved, (pos, vel, vol) = run_split(lambda: O.OracleMpm(dtype), OracleHaloBackend, oracle_step, dtype,
                                                    world, steps=60, every=4)
    assert moved > len(pos) // 10, moved                      # the test really migrates
    full = O.OracleMpm(dtype)
    full.AddParticles(pos, vel, vol, **MAT)
    full.Finalize()
    for it in range(60):
        full.substep(DT, -1, it % 7 == 0)
    ref = np.empty((len(pos), 3))
    ref[full.read("pids")] = full.read("positions").reshape(-1, 3)
    got = gathered(parts, migs, len(pos))
    assert np.max(np.abs(got - ref)) <= 1e-12
    # every particle sits in its slab (+ margin) and the bookkeeping is a partition of the scene
    gids = np.concatenate([m.gid[m.gid >= 0] for m in migs])
    assert np.array_equal(np.sort(gids), np.arange(len(pos)))
    for m, st in zip(migs, parts):
        lo, hi = m.bounds()
        x = got[m.gid[m.gid >= 0], 0]
        # migration ran 4 substeps ago at most: a particle moves < 0.2 cells per substep here
        assert np.all(x >= lo - DX) and np.all(x < hi + DX)


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dtype = np.float64
    pos, vel, vol = moving_block(dtype)
    edges = slab.slab_edges(pos, world)
    own = slab.owner_of(pos[:, 0], edges)
    st, gid = make_part(lambda: O.OracleMpm(dtype), dtype, pos, vel, vol, own == rank, rank, spare=len(pos) // 2)
    backend = OracleHaloBackend(st)
    ex = slab.SlabExchange(backend)
    mig = slab.SlabMigrator(st, rank, world, edges, gid, park_of(rank), margin=0.5 * DX, cap=len(pos),
                            all_gather=slab.dist_all_gather())
    moved = 0
    for it in range(40):
        st.RebuildMapping(it % 7 == 0); st.CalcFemStateAndForce(DT); st.ParticleToGrid(DT)
        ex.exchange()
        st.UpdateGrid(-1); st.GridToParticle(DT)
        if (it + 1) % 4 == 0:
            n = mig.migrate()
            moved += n
            if n:
                ex.planned = False
                backend.reset_plan()
    v = st.read("positions").reshape(-1, 3)
    by_pid = np.empty_like(v)
    by_pid[st.read("pids")] = v
    act = mig.gid >= 0
    ret[rank] = (mig.gid[act], by_pid[act], moved)
    dist.destroy_process_group()


def test_gloo_world2_slabs_with_migration_match_single_state():
    world, port = 2, free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        ps = [mp.Process(target=_gloo_worker, args=(r, world, port, ret)) for r in range(world)]
        [p.start() for p in ps]
        [p.join(180) for p in ps]
        assert all(p.exitcode == 0 for p in ps), [p.exitcode for p in ps]
        res = dict(ret)
    dtype = np.float64
    pos, vel, vol = moving_block(dtype)
    full = O.OracleMpm(dtype)
    full.AddParticles(pos, vel, vol, **MAT)
    full.Finalize()
    for it in range(40):
        full.substep(DT, -1, it % 7 == 0)
    ref = np.empty((len(pos), 3))
    ref[full.read("pids")] = full.read("positions").reshape(-1, 3)
    got = np.full((len(pos), 3), np.nan)
    for r in range(world):
        got[res[r][0]] = res[r][1]
    assert res[0][2] == res[1][2] > 100
    assert not np.isnan(got).any() and np.max(np.abs(got - ref)) <= 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol,world", [(np.float64, 1e-10, 2), (np.float64, 1e-10, 3), (np.float32, 5e-4, 2)])
def test_gpu_slabs_with_migration_match_single_state_and_oracle(dtype, tol, world):
    import torch
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    sv = GpuMpmSolver()
    parts_box = []

    def sync_all():
        for s in parts_box:
            sv.GpuSync(s)
        torch.cuda.synchronize()

    def make_state():
        s = GpuMpmState(dtype)
        parts_box.append(s)
        return s

    def step(parts, ex, it):
        for s in parts:
            sv.RebuildMapping(s, it % 7 == 0); sv.CalcFemStateAndForce(s, DT); sv.ParticleToGrid(s, DT)
        ex.exchange()
        for s in parts:
            sv.UpdateGrid(s, -1); sv.GridToParticle(s, DT)

    parts, migs, moved, (pos, vel, vol) = run_split(make_state, slab.GpuHaloBackend, step, dtype, world, steps=60,
                                                    every=4, sync=sync_all)
    assert moved > len(pos) // 10
    full, orc = GpuMpmState(dtype), O.OracleMpm(dtype)
    for s in (full, orc):
        s.AddParticles(pos, vel, vol, **MAT)
        s.Finalize()
    for it in range(60):
        s = it % 7 == 0
        sv.RebuildMapping(full, s); sv.CalcFemStateAndForce(full, DT); sv.ParticleToGrid(full, DT)
        sv.UpdateGrid(full, -1); sv.GridToParticle(full, DT)
        orc.substep(DT, -1, s)
    got = gathered(parts, migs, len(pos))
    for ref_state in (full, orc):
        ref = np.empty((len(pos), 3))
        ref[ref_state.read("pids")] = np.asarray(ref_state.read("positions"), np.float64).reshape(-1, 3)
        assert np.max(np.abs(got - ref)) <= tol * np.max(np.abs(ref))
    gids = np.concatenate([m.gid[m.gid >= 0] for m in migs])
    assert np.array_equal(np.sort(gids), np.arange(len(pos)))
    # the inactive particles stay parked and massless
    for st, m in zip(parts, migs):
        vol_by_pid = np.empty(st.n_particles())
        vol_by_pid[st.read("pids")] = st.read("volumes")
        assert np.all(vol_by_pid[m.gid < 0] == 0) and np.all(vol_by_pid[m.gid >= 0] > 0)
    for s in parts + [full]:
        s.Destroy()


PARK_FAR = [0.7, 0.7, 0.7, 0.9, 0.9, 0.9]
PARK_INSIDE_BODY = [0.41, 0.46, 0.31, 0.46, 0.49, 0.34]   # in the middle of moving_block(): the flow passes through it


@pytest.mark.gpu
@pytest.mark.parametrize("park", [PARK_FAR, PARK_INSIDE_BODY], ids=["far", "inside_body"])
def test_inactive_particles_are_invisible_to_the_grid(park):
    """Spare particles of volume 0 deposit nothing and do not mark grid blocks as touched (the exchange plans the
    reachable blocks from the mass on the grid): a state with spares steps exactly like one without, and its
    touched-block set is the same.  The spares do not follow the flow either: parked in the middle of the
    moving body they stay where they are, at rest."""
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    dtype = np.float64
    pos, vel, vol = moving_block(dtype)
    sv = GpuMpmSolver()
    outs = []
    for spare in (0, 5000):
        st = GpuMpmState(dtype)
        st.AddParticles(pos, vel, vol, **MAT)
        if spare:
            st.AddParticles(slab.park_points(park, spare), np.zeros((spare, 3)), 0.0, **MAT)
        st.Finalize()
        for it in range(8):
            sv.RebuildMapping(st, it == 3); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT)
            if it == 7:
                flags = st.read("grid_touched_flags").copy()
                mass = st.read("grid_masses").copy()
            sv.UpdateGrid(st, -1); sv.GridToParticle(st, DT)
        v = np.asarray(st.read("positions")).reshape(-1, 3)
        by_pid = np.empty_like(v)
        by_pid[st.read("pids")] = v
        w = np.asarray(st.read("velocities")).reshape(-1, 3)
        w_by_pid = np.empty_like(w)
        w_by_pid[st.read("pids")] = w
        outs.append((flags, mass, by_pid[:len(pos)], by_pid[len(pos):], w_by_pid[len(pos):]))
        st.Destroy()
    assert np.array_equal(outs[0][0], outs[1][0]) and outs[0][0].sum() > 0
    if park is PARK_FAR:
        assert np.array_equal(outs[0][1], outs[1][1])
        assert np.max(np.abs(outs[0][2] - outs[1][2])) <= 1e-13
    else:
        # spares that share grid blocks with the body add exact zeros, but they shift which particles share a chunk of
        # the block, i.e. the order of the sums: equal to rounding, not to the bit
        assert np.max(np.abs(outs[0][1] - outs[1][1])) <= 1e-12 * np.max(outs[0][1])
        assert np.max(np.abs(outs[0][2] - outs[1][2])) <= 1e-11
    assert np.array_equal(outs[1][3], slab.park_points(park, 5000))   # the spares never moved
    assert not outs[1][4].any()                                      # and are at rest


def test_oracle_inactive_particles_stay_parked_inside_a_moving_body():
    """The oracle's restatement of the same rule (G2P leaves an inactive spare where it is)."""
    dtype = np.float64
    pos, vel, vol = moving_block(dtype)
    st = O.OracleMpm(dtype)
    st.AddParticles(pos, vel, vol, **MAT)
    st.AddParticles(slab.park_points(PARK_INSIDE_BODY, 300), np.zeros((300, 3)), 0.0, **MAT)
    st.Finalize()
    for it in range(4):
        st.substep(DT, -1, it == 2)
    v = np.asarray(st.read("positions")).reshape(-1, 3)
    by_pid = np.empty_like(v)
    by_pid[st.read("pids")] = v
    assert np.array_equal(by_pid[len(pos):], slab.park_points(PARK_INSIDE_BODY, 300))
    assert np.max(np.abs(by_pid[:len(pos)] - pos)) > 1e-4   # the body itself moved


@pytest.mark.gpu
def test_migrate_in_refuses_a_material_index_outside_the_table():
    """A record's material index is an index into the RECEIVING state's table: one that the state never
    registered, or that is not an integer, must be refused on the host before any kernel indexes the table with it."""
    from drake_b200._lib import GmpmError
    from drake_b200.gpu_mpm import GpuMpmState
    dtype = np.float64
    pos, vel, vol = moving_block(dtype)
    st = GpuMpmState(dtype)
    st.AddParticles(pos, vel, vol, **MAT)
    st.AddParticles(slab.park_points(PARK_FAR, 64), np.zeros((64, 3)), 0.0, **MAT)
    st.Finalize()
    rec, pids = st.migrate_out(0, 0.0, 0.42, PARK_FAR, 64)
    assert len(pids) > 3 and np.all(rec[:, 25] == 0)
    for bad in (1.0, 16.0, 255.0, -1.0, 0.5, np.nan):
        r = rec.copy()
        r[2, 25] = bad
        with pytest.raises(GmpmError, match="material index") as e:
            st.migrate_in(r)
        assert e.value.status == -1
    st.migrate_in(rec)    # nothing was consumed by the refused calls: the spares are all still there
    vol_now = np.asarray(st.read("volumes"))
    assert int((vol_now > 0).sum()) == len(pos)
    st.Destroy()


@pytest.mark.gpu
def test_checkpoint_taken_after_a_migration_restores_the_material_of_every_slot():
    """Two materials in one state.  Particles of both leave and come back into each other's slots, so the slot ->
    material map no longer is what the AddParticles calls produce.  A checkpoint taken now and loaded into a state rebuilt by the same calls must continue like the
    uninterrupted run (version 2 of the blob dropped the material indices and silently applied the wrong
    law)."""
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    dtype = np.float64
    sv = GpuMpmSolver()
    pos, vel, vol = moving_block(dtype)
    soft = dict(model="fixed_corotated", youngs=2e4, poisson=0.3)
    stiff = dict(model="stvk_hencky", youngs=4e5, poisson=0.2)
    left = pos[:, 0] < 0.44

    def build():
        st = GpuMpmState(dtype)
        st.AddParticles(pos[left], vel[left], vol, **soft)
        st.AddParticles(pos[~left], vel[~left], vol, **stiff)
        st.AddParticles(slab.park_points(PARK_FAR, 2000), np.zeros((2000, 3)), 0.0, **soft)
        st.Finalize()
        return st

    def run(st, n, first):
        for it in range(first, first + n):
            sv.RebuildMapping(st, it % 5 == 0); sv.CalcFemStateAndForce(st, DT); sv.ParticleToGrid(st, DT)
            sv.UpdateGrid(st, -1); sv.GridToParticle(st, DT)

    def by_pid(st, name, comps):
        v = np.asarray(st.read(name), np.float64).reshape(-1, comps)
        out = np.empty_like(v)
        out[st.read("pids")] = v
        return out

    def swap_places(st, keep_material=True):
        """Soft particles at the rear and stiff ones at the front leave; the stiff records are seated first, so they
        take the lowest free slots -- the ones the soft leavers vacated -- and the soft records the stiff leavers'."""
        rec_soft, _ = st.migrate_out(0, 0.415, 1e30, PARK_FAR, 2000)
        rec_stiff, _ = st.migrate_out(0, -1e30, 0.47, PARK_FAR, 2000)
        assert 50 < len(rec_stiff) < 2000 and 50 < len(rec_soft) < 2000
        assert np.all(rec_soft[:, 25] == 0) and np.all(rec_stiff[:, 25] == 1)
        if not keep_material:     # what a restore without the material indices amounts to: the slot's registered law
            rec_soft[:, 25], rec_stiff[:, 25] = 1, 0
        pids_stiff = st.migrate_in(rec_stiff)
        pids_soft = st.migrate_in(rec_soft)
        # stiff particles now live in slots registered as soft, and the other way round
        assert np.sum(pids_stiff < left.sum()) == min(len(rec_stiff), len(rec_soft))
        assert np.any((pids_soft >= left.sum()) & (pids_soft < len(pos)))

    a = build()
    run(a, 6, 0)
    swap_places(a)
    blob = a.save_state()
    run(a, 12, 6)
    b = build()
    b.load_state(blob)
    run(b, 12, 6)
    for name, comps in (("positions", 3), ("velocities", 3), ("particle_F", 9)):
        if name == "particle_F":
            va, vb = np.asarray(a.read(name)).reshape(-1, 9), np.asarray(b.read(name)).reshape(-1, 9)
        else:
            va, vb = by_pid(a, name, comps), by_pid(b, name, comps)
        assert np.max(np.abs(va - vb)) <= 1e-10 * np.max(np.abs(va)), name
    # control: the two laws do part ways on this motion -- a state that keeps the material table of the AddParticles
    # calls (the arrivals' slots registered as soft) ends up measurably elsewhere
    c = build()
    run(c, 6, 0)
    swap_places(c, keep_material=False)
    run(c, 12, 6)
    assert np.max(np.abs(by_pid(c, "positions", 3) - by_pid(a, "positions", 3))) > 1e-6
    for s in (a, b, c):
        s.Destroy()


def test_packed_transport_path_of_the_migrator_on_two_threads():
    """SlabMigrator.migrate() with an all_gather that carries one packed array per rank (the NCCL tensor path of
    slab.dist_all_gather_packed), driven by two threads standing in for two ranks: same result as the object path."""
    import threading
    dtype = np.float64
    world = 2
    pos, vel, vol = moving_block(dtype)
    edges = slab.slab_edges(pos, world)
    own = slab.owner_of(pos[:, 0], edges)
    barrier = threading.Barrier(world)
    box = [None] * world

    def make_gather(rank):
        def gather(rows):
            box[rank] = rows
            barrier.wait()
            out = np.concatenate([b for b in box if len(b)]) if any(len(b) for b in box) else np.zeros((0, 28))
            barrier.wait()
            return out
        gather.packed = True
        return gather

    parts, migs = [], []
    for r in range(world):
        st, gid = make_part(lambda: O.OracleMpm(dtype), dtype, pos, vel, vol, own == r, r, spare=len(pos) // 2)
        parts.append(st)
        migs.append(slab.SlabMigrator(st, r, world, edges, gid, park_of(r), margin=0.5 * DX, cap=len(pos),
                                      all_gather=make_gather(r)))
    ex = slab.InProcessExchange([OracleHaloBackend(s) for s in parts])
    moved = [0] * world
    for it in range(24):
        oracle_step(parts, ex, it)
        if (it + 1) % 4 == 0:
            def work(r):
                moved[r] += migs[r].migrate()
            ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            ex.planned = False
            for b in ex.backends:
                b.reset_plan()
    assert moved[0] == moved[1] > 50
    full = O.OracleMpm(dtype)
    full.AddParticles(pos, vel, vol, **MAT)
    full.Finalize()
    for it in range(24):
        full.substep(DT, -1, it % 7 == 0)
    ref = np.empty((len(pos), 3))
    ref[full.read("pids")] = full.read("positions").reshape(-1, 3)
    assert np.max(np.abs(gathered(parts, migs, len(pos)) - ref)) <= 1e-12


# ------------------------------------------------------------------------------------------------------------------
# a migration that cannot be completed must be refused before anything is taken out of a state
# ------------------------------------------------------------------------------------------------------------------
def _short_of_room(make_state, make_backend, step, dtype, sync=None):
    """Two slabs, the right one with 3 spare slots only; stepped until more than 3 particles head for it."""
    pos, vel, vol = moving_block(dtype)
    edges = slab.slab_edges(pos, 2)
    own = slab.owner_of(pos[:, 0], edges)
    parts, migs = [], []
    for r in range(2):
        st, gid = make_part(make_state, dtype, pos, vel, vol, own == r, r, spare=(len(pos) // 2 if r == 0 else 3))
        parts.append(st)
        migs.append(slab.SlabMigrator(st, r, 2, edges, gid, park_of(r), margin=0.5 * DX, cap=4))
    ex = slab.InProcessExchange([make_backend(s) for s in parts], sync=sync)
    for it in range(24):
        step(parts, ex, it)
    return parts, migs, ex, edges


def _snapshot(parts):
    return [(np.array(st.read("positions")), np.array(st.read("volumes")), np.array(st.read("pids"))) for st in parts]


def test_migration_without_room_is_refused_on_every_rank_and_moves_nothing():
    dtype = np.float64
    parts, migs, ex, edges = _short_of_room(lambda: O.OracleMpm(dtype), OracleHaloBackend, oracle_step, dtype)
    rows = []
    for m in migs:
        leavers, free = m.state.migrate_count(m.axis, *m.bounds(), m.edges)
        rows.append(np.concatenate([[m.rank, free], leavers]).astype(np.float64)[None, :])
    assert rows[0][0, 3] > 3 and rows[1][0, 1] == 3          # more than 3 head right, 3 seats there
    assert rows[0][0, 3] > migs[0].cap                       # and more than one capped migrate_out call would report
    before = _snapshot(parts)
    with pytest.raises(slab.MigrationOverflow):
        slab.migrate_in_process(migs, ex)
    # the collective form: every rank sees the same table and raises on its own
    for m in migs:
        m.all_gather = lambda row: [r.copy() for r in rows]
        with pytest.raises(slab.MigrationOverflow, match="slab 1 would receive"):
            m.migrate()
    after = _snapshot(parts)
    for b, a in zip(before, after):
        assert all(np.array_equal(x, y) for x, y in zip(b, a))
    assert all(np.array_equal(m.gid >= 0, np.arange(len(m.gid)) < (m.gid >= 0).sum()) for m in migs)   # nobody was parked


@pytest.mark.gpu
def test_gpu_migrate_count_is_a_dry_run_and_agrees_with_oracle_and_with_migrate_out():
    import torch
    from drake_b200.gpu_mpm import GpuMpmSolver, GpuMpmState
    sv = GpuMpmSolver()
    for dtype in (np.float64, np.float32):
        boxes = []

        def make_state():
            s = GpuMpmState(dtype)
            boxes.append(s)
            return s

        def step(parts, ex, it):
            for s in parts:
                sv.RebuildMapping(s, it % 7 == 0); sv.CalcFemStateAndForce(s, DT); sv.ParticleToGrid(s, DT)
            ex.exchange()
            for s in parts:
                sv.UpdateGrid(s, -1); sv.GridToParticle(s, DT)

        def sync_all():
            for s in boxes:
                sv.GpuSync(s)
            torch.cuda.synchronize()

        parts, migs, ex, edges = _short_of_room(make_state, slab.GpuHaloBackend, step, dtype, sync=sync_all)
        oparts, omigs, _, _ = _short_of_room(lambda: O.OracleMpm(dtype), OracleHaloBackend, oracle_step, dtype)
        before = _snapshot(parts)
        for m, om in zip(migs, omigs):
            got = m.state.migrate_count(m.axis, *m.bounds(), m.edges)
            want = om.state.migrate_count(om.axis, *om.bounds(), om.edges)
            assert np.array_equal(got[0], want[0]) and got[1] == want[1], (dtype, got, want)
        leavers0 = int(migs[0].state.migrate_count(0, *migs[0].bounds(), edges)[0].sum())
        assert leavers0 > 4
        with pytest.raises(slab.MigrationOverflow):
            slab.migrate_in_process(migs, ex)
        after = _snapshot(parts)
        for b, a in zip(before, after):
            assert all(np.array_equal(x, y) for x, y in zip(b, a))
        # the count is what an uncapped migrate_out takes (a capped one, cap = 4, reports 4 and hides the rest)
        rec, pids = parts[0].migrate_out(0, *migs[0].bounds(), park_of(0), 4)
        assert len(pids) == 4
        rec2, pids2 = parts[0].migrate_out(0, *migs[0].bounds(), park_of(0), 1 << 16)
        assert len(pids) + len(pids2) == leavers0
        for s in parts:
            s.Destroy()
