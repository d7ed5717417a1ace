"""Multi-GPU path: host-side protocol on CPU (gloo, world_size 2, oracle backend) and the device side
on one GPU (two states in one process) and on two GPUs (NCCL).  The pinned property everywhere:
a scene split over R ranks gives the same particles as the same scene on one rank."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle_py as O  # noqa: E402
from drake_b200 import scenes, slab  # noqa: E402
from oracle_halo import OracleHaloBackend, oracle_substep  # noqa: E402

DT = 1e-3


def sheets_scene(dtype, n_sheets=4, res=24):
    """Warm wavy sheets 2.7 dx apart: neighbouring sheets reach common grid blocks."""
    rng = np.random.default_rng(11)
    out = []
    for s in range(n_sheets):
        pos, vel, tri = scenes.cloth_sheet(res, 0.2, 0.33 + 0.011 * s, 0.35, 0.45 + 0.021 * s, dtype)
        pos = pos.astype(np.float64)
        pos[:, 2] += 0.006 * np.sin(9 * pos[:, 0]) * np.cos(7 * pos[:, 1])
        out.append((pos.astype(dtype), rng.normal(0, 0.3, pos.shape).astype(dtype), tri))
    return out


def make_oracle(dtype, sheets, **params):
    orc = O.OracleMpm(dtype, O.default_params(**params))
    for pos, vel, tri in sheets:
        orc.AddQRCloth(pos, vel, tri)
    orc.Finalize()
    return orc


def vert_range(sheets, part):
    n = [s[0].shape[0] for s in sheets]
    lo = sum(n[:part.start])
    return lo, lo + sum(n[part.start:part.stop])


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_partition_meshes():
    assert [list(r) for r in slab.partition_meshes(5, 2)] == [[0, 1, 2], [3, 4]]
    assert [len(r) for r in slab.partition_meshes(34 * 8, 8)] == [34] * 8
    assert [list(r) for r in slab.partition_meshes(2, 3)] == [[0], [1], []]


def test_oracle_two_parts_in_process_match_single():
    dtype = np.float64
    sheets = sheets_scene(dtype)
    full = make_oracle(dtype, sheets)
    parts = slab.partition_meshes(len(sheets), 2)
    orcs = [make_oracle(dtype, [sheets[i] for i in p]) for p in parts]
    ex = slab.InProcessExchange([OracleHaloBackend(o) for o in orcs])
    shared_seen = 0
    for it in range(8):
        full.substep(DT, -1, it == 3)
        for o in orcs:
            o.RebuildMapping(it == 3)
            o.CalcFemStateAndForce(DT)
            o.ParticleToGrid(DT)
        shared_seen += ex.exchange()
        for o in orcs:
            o.UpdateGrid(-1)
            o.GridToParticle(DT)
    assert shared_seen > 0
    ref, _ = full.DumpCpuState()
    for o, p in zip(orcs, parts):
        lo, hi = vert_range(sheets, p)
        got, _ = o.DumpCpuState()
        assert np.max(np.abs(got - ref[lo:hi])) <= 1e-12 * np.max(np.abs(ref))


def _gloo_worker(rank, world, port, ret):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        dtype = np.float64
        sheets = sheets_scene(dtype)
        part = slab.partition_meshes(len(sheets), world)[rank]
        orc = make_oracle(dtype, [sheets[i] for i in part])
        ex = slab.SlabExchange(OracleHaloBackend(orc))
        shared = 0
        for it in range(8):
            oracle_substep(orc, lambda: ex.exchange(), DT, -1, it == 3)
            shared += ex.last_n_shared
        got, _ = orc.DumpCpuState()
        full = make_oracle(dtype, sheets)
        for it in range(8):
            full.substep(DT, -1, it == 3)
        ref, _ = full.DumpCpuState()
        lo, hi = vert_range(sheets, part)
        ret[rank] = (float(np.max(np.abs(got - ref[lo:hi])) / np.max(np.abs(ref))), shared, ex.bytes_reduced)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_oracle_backend_matches_single():
    import torch.multiprocessing as mp
    world, port = 2, free_port()
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_gloo_worker, args=(world, port, ret), nprocs=world, join=True)
        assert len(ret) == world
        for r in range(world):
            err, shared, nbytes = ret[r]
            assert err <= 1e-12, (r, err)
            assert shared > 0 and nbytes > 0
        assert ret[0][1] == ret[1][1]          # every rank planned the same shared-block lists


def test_partition_by_size_minimises_the_fullest_rank():
    import itertools
    rng = np.random.default_rng(5)
    for n, world in [(1, 1), (3, 2), (7, 3), (9, 4), (5, 8), (12, 5)]:
        sizes = rng.integers(1, 50, size=n)
        parts = slab.partition_meshes_by_size(sizes, world)
        assert len(parts) == world
        assert [i for p in parts for i in p] == list(range(n))          # contiguous cover, in order
        got = max((sum(sizes[i] for i in p) for p in parts), default=0)
        k = min(world, n)
        brute = min(max(sum(sizes[a:b]) for a, b in zip((0,) + c, c + (n,)))
                    for c in itertools.combinations(range(1, n), k - 1)) if k > 1 else sum(sizes)
        assert got == brute, (n, world, got, brute)
    assert slab.partition_meshes_by_size([], 3) == [range(0, 0)] * 3
    # equal sizes: the same load balance as the plain split
    eq = slab.partition_meshes_by_size([10] * 8, 4)
    assert sorted(len(p) for p in eq) == [2, 2, 2, 2]


@pytest.mark.parametrize("n_ranks", [2, 3])
def test_peer_window_protocol_on_the_oracle(n_ranks):
    """slab.InProcessExchange(peer=True) over oracle backends with numpy windows: sharer-set flags, sends into the
    sharers' windows, rank-ordered sums.  A scene split over the ranks reproduces the single-oracle run and the
    ranks hold identical bits on the blocks they share (3 ranks: some blocks have three sharers)."""
    dtype = np.float64
    sheets = sheets_scene(dtype, n_sheets=6, res=16)
    if n_ranks == 3:     # 1.2 dx apart instead of 2.7: the outer ranks' stencils meet inside the middle rank's blocks
        sheets = [(p - np.array([0.0, 0.0, 0.012 * s]), v, t) for s, (p, v, t) in enumerate(sheets)]
    full = make_oracle(dtype, sheets)
    parts = slab.partition_meshes(len(sheets), n_ranks)
    orcs = [make_oracle(dtype, [sheets[i] for i in p]) for p in parts]
    backs = [OracleHaloBackend(o) for o in orcs]
    ex = slab.InProcessExchange(backs, peer=True, peer_cap_blocks=4096)
    shared = 0
    for it in range(8):
        sort = it == 3
        full.substep(DT, -1, sort)
        for o in orcs:
            o.RebuildMapping(sort)
            o.CalcFemStateAndForce(DT)
            o.ParticleToGrid(DT)
        shared += ex.exchange()
        if it == 7:      # identical bits on every pair of sharers, checked on the exchanged sums
            for a in range(n_ranks):
                for b in range(a + 1, n_ranks):
                    both = np.repeat(backs[a].own & backs[b].own, 64)
                    ma, mb = orcs[a].view("grid_masses"), orcs[b].view("grid_masses")
                    assert np.array_equal(ma[both], mb[both])
                    pa, pb = orcs[a].view("grid_momentum").reshape(-1, 3), orcs[b].view("grid_momentum").reshape(-1, 3)
                    assert np.array_equal(pa[both], pb[both])
        for o in orcs:
            o.UpdateGrid(-1)
            o.GridToParticle(DT)
    assert shared > 0
    if n_ranks == 3:
        assert np.any(np.array([bin(int(s)).count("1") for s in backs[0].sharers]) == 3)
    ref, _ = full.DumpCpuState()
    for o, p in zip(orcs, parts):
        lo, hi = vert_range(sheets, p)
        got, _ = o.DumpCpuState()
        assert np.max(np.abs(got - ref[lo:hi])) <= 1e-12 * np.max(np.abs(ref))


# ---- device side ----------------------------------------------------------------------------------
def make_state(dtype, sheets, **params):
    from drake_b200.gpu_mpm import GpuMpmState
    st = GpuMpmState(dtype, **params)
    for pos, vel, tri in sheets:
        st.AddQRCloth(pos, vel, tri)
    st.Finalize()
    return st


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-10), (np.float32, 2e-4)])
def test_two_states_one_gpu_match_single_state_and_oracle(dtype, tol):
    from drake_b200.gpu_mpm import GpuMpmSolver
    sheets = sheets_scene(dtype)
    sv = GpuMpmSolver()
    full = make_state(dtype, sheets)
    orc = make_oracle(dtype, sheets)
    parts = slab.partition_meshes(len(sheets), 2)
    sts = [make_state(dtype, [sheets[i] for i in p]) for p in parts]
    def sync_all():
        for s in sts:
            sv.GpuSync(s)
        import torch
        torch.cuda.synchronize()

    ex = slab.InProcessExchange([slab.GpuHaloBackend(s) for s in sts], sync=sync_all)
    shared_seen = 0
    for it in range(12):
        sort = it == 5
        slab.substep(sv, full, None, DT, -1, sort)
        orc.substep(DT, -1, sort)
        for s in sts:
            sv.RebuildMapping(s, sort)
            sv.CalcFemStateAndForce(s, DT)
            sv.ParticleToGrid(s, DT)
        shared_seen += ex.exchange()
        for s in sts:
            sv.UpdateGrid(s, -1)
            sv.GridToParticle(s, DT)
    assert shared_seen > 0
    ref, _ = full.DumpCpuState()
    oref, _ = orc.DumpCpuState()
    scale = np.max(np.abs(oref))
    assert np.max(np.abs(ref - oref)) <= tol * scale
    for s, p in zip(sts, parts):
        lo, hi = vert_range(sheets, p)
        got, _ = s.DumpCpuState()
        assert np.max(np.abs(got - ref[lo:hi])) <= tol * scale
        assert np.max(np.abs(got - oref[lo:hi])) <= tol * scale


@pytest.mark.gpu
def test_two_states_one_gpu_overlapped_parts_match_plain_order():
    """UpdateGrid / GridToParticle split around the exchange (interior first, exchanged blocks after the
    sums are back) gives the same particles as the plain order; alone, part 0 is everything."""
    from drake_b200.gpu_mpm import GpuMpmSolver
    import torch
    dtype = np.float64
    sheets = sheets_scene(dtype)
    sv = GpuMpmSolver()
    parts = slab.partition_meshes(len(sheets), 2)

    def run(overlap):
        sts = [make_state(dtype, [sheets[i] for i in p]) for p in parts]

        def sync_all():
            for s in sts:
                sv.GpuSync(s)
            torch.cuda.synchronize()

        ex = slab.InProcessExchange([slab.GpuHaloBackend(s) for s in sts], sync=sync_all)
        shared = 0
        for it in range(10):
            for s in sts:
                sv.RebuildMapping(s, it == 4)
                sv.CalcFemStateAndForce(s, DT)
                sv.ParticleToGrid(s, DT)
            if overlap:
                shared += ex.begin()
                for s in sts:
                    s.update_grid_part(-1, 0)
                    s.g2p_part(DT, 0)
                ex.finish()
                for s in sts:
                    s.update_grid_part(-1, 1)
                    s.g2p_part(DT, 1)
            else:
                shared += ex.exchange()
                for s in sts:
                    sv.UpdateGrid(s, -1)
                    sv.GridToParticle(s, DT)
        assert shared > 0
        return [(s.read("positions"), s.read("velocities"), s.read("affine_matrices")) for s in sts]

    a, b = run(False), run(True)       # (vertex forces are summed with atomics: equal to rounding, not bitwise)
    for ra, rb in zip(a, b):
        for x, y in zip(ra, rb):
            assert np.max(np.abs(x - y)) <= 1e-9 * max(1.0, np.max(np.abs(x)))
    # a state that exchanges nothing: the two parts together are the plain calls
    st1, st2 = make_state(dtype, sheets), make_state(dtype, sheets)
    for it in range(3):
        for st in (st1, st2):
            sv.RebuildMapping(st, False)
            sv.CalcFemStateAndForce(st, DT)
            sv.ParticleToGrid(st, DT)
        sv.UpdateGrid(st1, -1)
        sv.GridToParticle(st1, DT)
        st2.update_grid_part(-1, 0)
        st2.g2p_part(DT, 0)
        st2.update_grid_part(-1, 1)
        st2.g2p_part(DT, 1)
    assert np.max(np.abs(st1.read("positions") - st2.read("positions"))) <= 1e-13
    c1, c2 = st1.read("affine_matrices"), st2.read("affine_matrices")
    assert np.max(np.abs(c1 - c2)) <= 1e-9 * max(1.0, np.max(np.abs(c1)))


@pytest.mark.gpu
@pytest.mark.parametrize("n_states,overlap", [(2, False), (3, False), (2, True)])
def test_peer_window_exchange_matches_single_state(n_states, overlap):
    """Peer mode: the block sums travel through receive windows (stores into the sharers' windows, delivery
    flags, device-side wait, rank-ordered sums) instead of an all-reduce.  Several states on ONE GPU share
    their windows by raw pointer; with 3 states some blocks have three sharers."""
    from drake_b200.gpu_mpm import GpuMpmSolver
    import torch
    dtype = np.float64
    sheets = sheets_scene(dtype, n_sheets=6, res=40)
    if n_states == 3:    # 1.2 dx apart: the outer states' stencils meet inside the middle state's blocks
        sheets = [(p - np.array([0.0, 0.0, 0.012 * s]), v, t) for s, (p, v, t) in enumerate(sheets)]
    sv = GpuMpmSolver()
    full = make_state(dtype, sheets)
    parts = slab.partition_meshes(len(sheets), n_states)
    sts = [make_state(dtype, [sheets[i] for i in p]) for p in parts]

    def sync_all():
        for s in sts:
            sv.GpuSync(s)
        torch.cuda.synchronize()

    ex = slab.InProcessExchange([slab.GpuHaloBackend(s) for s in sts], sync=sync_all, peer=True, peer_cap_blocks=4096)
    shared = 0
    for it in range(12):
        sort = it == 5
        slab.substep(sv, full, None, DT, -1, sort)
        for s in sts:
            sv.RebuildMapping(s, sort)
            sv.CalcFemStateAndForce(s, DT)
            sv.ParticleToGrid(s, DT)
        if overlap:
            shared += ex.begin()
            for s in sts:
                s.update_grid_part(-1, 0)
                s.g2p_part(DT, 0)
            ex.finish()
            for s in sts:
                s.update_grid_part(-1, 1)
                s.g2p_part(DT, 1)
        else:
            shared += ex.exchange()
            for s in sts:
                sv.UpdateGrid(s, -1)
                sv.GridToParticle(s, DT)
    assert shared > 0
    ref, _ = full.DumpCpuState()
    scale = np.max(np.abs(ref))
    for s, p in zip(sts, parts):
        lo, hi = vert_range(sheets, p)
        got, _ = s.DumpCpuState()
        assert np.max(np.abs(got - ref[lo:hi])) <= 1e-10 * scale
    # shared blocks carry identical bits on every sharer: grid velocities of the last substep agree exactly
    v = [s.read("grid_v_star") for s in sts]
    m = [s.read("grid_masses") for s in sts]
    tf = [s.read("grid_touched_flags") for s in sts]
    both = np.repeat((tf[0] > 0) & (tf[1] > 0), 64)    # cells of the blocks both ranks touched (block-major keys)
    assert both.any()
    np.testing.assert_array_equal(m[0][both], m[1][both])
    np.testing.assert_array_equal(v[0].reshape(-1, 3)[both], v[1].reshape(-1, 3)[both])
    if n_states == 3:                          # blocks all three states touched: identical bits on all of them
        all3 = np.repeat((tf[0] > 0) & (tf[1] > 0) & (tf[2] > 0), 64)
        assert all3.any()
        np.testing.assert_array_equal(m[0][all3], m[2][all3])
        np.testing.assert_array_equal(v[1].reshape(-1, 3)[all3], v[2].reshape(-1, 3)[all3])
    with pytest.raises(Exception):
        sts[0].halo_pack()                     # a peer-mode state has no packed buffer to all-reduce


@pytest.mark.gpu
def test_halo_call_order_is_checked():
    from drake_b200.gpu_mpm import GmpmError, GpuMpmSolver
    st = make_state(np.float64, sheets_scene(np.float64, n_sheets=1))
    sv = GpuMpmSolver()
    with pytest.raises(GmpmError):
        st.halo_flags()                        # no P2G yet
    sv.RebuildMapping(st, False)
    sv.CalcFemStateAndForce(st, DT)
    sv.ParticleToGrid(st, DT)
    with pytest.raises(GmpmError):
        st.halo_unpack()                       # nothing packed
    with pytest.raises(GmpmError):
        st.halo_pack()                         # nothing planned
    ptr, n = st.halo_flags()
    assert ptr and n == 32768
    st.halo_plan_async()
    n_shared, _ = st.halo_pack()               # alone: no block has flags > 1
    assert n_shared == 0
    st.halo_unpack()
    sv.UpdateGrid(st, -1)
    sv.GridToParticle(st, DT)


def _nccl_worker(rank, world, port, ret, peer=False):
    import torch
    import torch.distributed as dist
    from drake_b200.gpu_mpm import GpuMpmSolver
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        dtype = np.float64
        sheets = sheets_scene(dtype, n_sheets=6, res=40)
        part = slab.partition_meshes(len(sheets), world)[rank]
        st = make_state(dtype, [sheets[i] for i in part], device=rank)
        sv = GpuMpmSolver()
        ex = slab.SlabExchange(slab.GpuHaloBackend(st), peer=peer, peer_cap_blocks=4096)
        shared = 0
        for it in range(20):
            slab.substep(sv, st, ex, DT, -1, it == 7, overlap=(it % 2 == 1))   # both orders
            shared += ex.last_n_shared
        got, _ = st.DumpCpuState()
        full = make_oracle(dtype, sheets)
        for it in range(20):
            full.substep(DT, -1, it == 7)
        ref, _ = full.DumpCpuState()
        lo, hi = vert_range(sheets, part)
        ret[rank] = (float(np.max(np.abs(got - ref[lo:hi])) / np.max(np.abs(ref))), shared)
        st.Destroy()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("peer", [False, True])
def test_nccl_world2_matches_single_process_oracle(peer):
    """peer=True: windows mapped across the two processes through CUDA IPC, stores over NVLink."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world, port = 2, free_port()
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_nccl_worker, args=(world, port, ret, peer), nprocs=world, join=True)
        for r in range(world):
            err, shared = ret[r]
            assert err <= 1e-9, (r, err)
            assert shared > 0
