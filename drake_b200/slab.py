"""Multi-GPU driver of the MPM hot path: one process and one GpuMpmState per GPU, scenes partitioned by
whole cloth meshes, grid blocks reached by several ranks summed between ParticleToGrid and UpdateGrid.

The reference is single-GPU (multibody/gpu_mpm/settings.h:40: G_DEVICE_COUNT = 1, never read); this is
the extension SURVEY.md 8(e) plans.  The device side lives in libgmpm_b200.so (csrc/halo.cuh, entry
points gmpm_halo_flags / halo_plan_async / halo_pack / halo_unpack); this module is the host plumbing:
it owns no arithmetic, it only orders the two collectives per substep between those calls.

    exchange (critical path, between P2G and UpdateGrid; list planned one substep earlier):
        packed = backend.halo_pack()       # T[n_shared * 256]: own mass + momentum of the shared blocks
        all_reduce(packed, SUM)            # NCCL over NVLink (gloo in the CPU tests)
        backend.halo_unpack()              # sums written back where this rank touched the block
    plan for the next substep (side stream, overlaps UpdateGrid / G2P / FEM / binning / P2G):
        flags = backend.halo_flags()       # uint8[n_blocks]: 1 within one block of a touched block
        all_reduce(flags, SUM)
        backend.halo_plan_async()          # shared = flags > 1, ascending ids, same list on every rank

A particle moves less than one grid block per substep, so the dilated set planned at substep k covers
everything substep k+1 can touch (checked on the device; halo_pack raises otherwise).  Nothing in the
substep blocks the host: the count of shared blocks travels through pinned memory one substep ahead.

Peer mode (`SlabExchange(..., peer=True)`, <= 8 ranks of one NVSwitch domain) takes the collective off the
critical path altogether: every rank owns a receive window in device memory that the others map through
CUDA IPC; `halo_peer_send` stores a rank's block values straight into the windows of the ranks sharing
each block and raises their delivery flags, `halo_peer_recv` waits for its sharers on the device and sums
in ascending rank order.  NCCL then only carries the 1-byte-per-block flags of the plan, on the side stream.

`backend` is anything with those four methods returning torch tensors that alias its buffers: the GPU
backend below wraps a GpuMpmState (tensors alias device memory, collectives are enqueued on the state's
own CUDA streams); the CPU tests drive the same protocol class with a backend built on the oracle.
"""
import numpy as np


def partition_meshes(n_meshes, world):
    """Contiguous split of mesh indices 0..n_meshes-1 over `world` ranks (sizes differ by at most one).
    Contiguous in stacking order = spatial slabs for stacked sheets."""
    base, extra = divmod(n_meshes, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append(range(start, start + n))
        start += n
    return out


def partition_meshes_by_size(sizes, world):
    """Contiguous split of meshes with UNEQUAL particle counts `sizes` over `world` ranks that minimises the
    largest rank (the substep time of the job is that of its fullest rank): the classic linear-partition
    dynamic programme, O(world * n^2).  Contiguity keeps the ranks spatial slabs for stacked scenes, which is
    what keeps the shared-block set small.  Returns a list of `range`s like partition_meshes; ranks beyond the
    number of meshes get empty ranges."""
    n = len(sizes)
    if n == 0 or world <= 0:
        return [range(0, 0) for _ in range(max(world, 0))]
    k = min(world, n)
    pre = np.concatenate([[0], np.cumsum(np.asarray(sizes, dtype=np.int64))])
    inf = np.iinfo(np.int64).max
    best = np.full((k + 1, n + 1), inf, dtype=np.int64)   # best[j][i]: min over splits of the first i meshes into j parts
    cut = np.zeros((k + 1, n + 1), dtype=np.int64)
    best[0][0] = 0
    for j in range(1, k + 1):
        for i in range(j, n + 1):
            for s in range(j - 1, i):
                if best[j - 1][s] == inf:
                    continue
                cost = max(int(best[j - 1][s]), int(pre[i] - pre[s]))
                if cost < best[j][i]:
                    best[j][i], cut[j][i] = cost, s
    bounds, i = [n], n
    for j in range(k, 0, -1):
        i = int(cut[j][i])
        bounds.append(i)
    bounds.reverse()
    out = [range(bounds[r], bounds[r + 1]) for r in range(k)]
    return out + [range(n, n) for _ in range(world - k)]


class _DevArray:
    """Zero-copy view of a raw device pointer for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


class GpuHaloBackend:
    """The four device steps of one GpuMpmState as torch tensors aliasing the library's buffers."""

    def __init__(self, state):
        import torch
        self.state = state
        self.torch = torch
        self.device = torch.device("cuda", state.params.device if state.params.device >= 0 else torch.cuda.current_device())
        self.stream = torch.cuda.ExternalStream(state.stream_handle(), device=self.device)
        self.plan_stream = torch.cuda.ExternalStream(state.halo_stream_handle(), device=self.device)
        self._typestr = "<f8" if state.dtype == np.float64 else "<f4"
        self._views = {}

    def _view(self, ptr, n, typestr):
        key = (ptr, n, typestr)
        t = self._views.get(key)
        if t is None:
            if len(self._views) > 64:
                self._views.clear()
            t = self._views[key] = self.torch.as_tensor(_DevArray(ptr, n, typestr), device=self.device)
        return t

    def halo_flags(self):
        ptr, n = self.state.halo_flags()
        return self._view(ptr, n, "|u1")

    def halo_plan_async(self):
        self.state.halo_plan_async()

    def reset_plan(self):
        self.state.halo_reset_plan()

    def halo_pack(self):
        n_shared, ptr = self.state.halo_pack()
        if n_shared == 0:
            return None
        return self._view(ptr, n_shared * 256, self._typestr)

    def halo_unpack(self):
        self.state.halo_unpack()

    def halo_peer_send(self):
        return self.state.halo_peer_send()

    def halo_peer_recv(self):
        self.state.halo_peer_recv()


class SlabExchange:
    """Runs the exchange protocol over a torch.distributed process group (NCCL on GPUs, gloo on CPU)."""

    def __init__(self, backend, group=None, peer=False, peer_cap_blocks=8192):
        import torch.distributed as dist
        self.backend = backend
        self.dist = dist
        self.group = group
        self.peer = bool(peer)
        if self.peer:
            # windows: allocate, publish the CUDA IPC handles, map everybody else's (once; they never move)
            rank, world = dist.get_rank(group), dist.get_world_size(group)
            st = backend.state
            st.halo_peer_init(rank, world, peer_cap_blocks)
            handle, _ = st.halo_peer_window()
            handles = [None] * world
            dist.all_gather_object(handles, handle, group=group)
            for q, h in enumerate(handles):
                if q != rank:
                    st.halo_peer_attach(q, ipc_handle=h)
            dist.barrier(group=group)
        self.stream = getattr(backend, "stream", None)
        self.plan_stream = getattr(backend, "plan_stream", None)
        self.planned = False
        self.last_n_shared = 0
        self.bytes_reduced = 0
        self._work = None

    def _all_reduce(self, t, stream):
        if stream is not None:
            import torch
            with torch.cuda.stream(stream):        # ordered with the state's kernels on that stream
                self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        else:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        self.bytes_reduced += t.numel() * t.element_size()

    def _plan(self):
        flags = self.backend.halo_flags()
        self._all_reduce(flags, self.plan_stream)
        self.backend.halo_plan_async()
        self.planned = True

    def begin(self):
        """After ParticleToGrid: pack this rank's share of the shared blocks and START their all-reduce.
        Work that does not depend on other ranks (update_grid_part(0), g2p_part(0)) goes between begin()
        and finish() and runs while the collective is in flight."""
        if not self.planned:
            self._plan()                          # start-up: plan from the current (dilated) touched set
        if self.peer:
            self.last_n_shared = self.backend.halo_peer_send()
            return self.last_n_shared
        packed = self.backend.halo_pack()         # every rank holds the same shared-block list
        self.last_n_shared = 0 if packed is None else packed.numel() // 256
        self._work = None
        if packed is not None:
            if self.stream is not None:
                import torch
                with torch.cuda.stream(self.stream):
                    self._work = self.dist.all_reduce(packed, op=self.dist.ReduceOp.SUM, group=self.group, async_op=True)
            else:
                self._work = self.dist.all_reduce(packed, op=self.dist.ReduceOp.SUM, group=self.group, async_op=True)
            self.bytes_reduced += packed.numel() * packed.element_size()
        return self.last_n_shared

    def finish(self):
        """Orders the state's stream after the all-reduce, writes the sums back, plans the next exchange."""
        if self.peer:
            self.backend.halo_peer_recv()
            self._plan()
            return
        if self._work is not None:
            if self.stream is not None:
                import torch
                with torch.cuda.stream(self.stream):
                    self._work.wait()             # stream-ordered: the host does not block
            else:
                self._work.wait()
            self._work = None
        self.backend.halo_unpack()
        self._plan()                              # for the next substep, off the critical path

    def exchange(self):
        """Call between ParticleToGrid and UpdateGrid, on every rank (no overlap: begin + finish)."""
        n = self.begin()
        self.finish()
        return n


class InProcessExchange:
    """The same protocol for several states living in ONE process (tests: two states on one GPU, or two
    oracles on the CPU): the all-reduces become explicit sums over the backends' tensors."""

    def __init__(self, backends, sync=None, peer=False, peer_cap_blocks=8192):
        self.backends = backends
        self.sync = sync or (lambda: None)        # GPU states use several streams: order them by hand
        self.planned = False
        self.peer = bool(peer)
        if self.peer:                             # same process: the windows are shared by raw pointer
            states = [b.state for b in backends]
            for r, st in enumerate(states):
                st.halo_peer_init(r, len(states), peer_cap_blocks)
            ptrs = [st.halo_peer_window()[1] for st in states]
            for r, st in enumerate(states):
                for q, p in enumerate(ptrs):
                    if q != r:
                        st.halo_peer_attach(q, pointer=p)

    @staticmethod
    def _sum(ts):
        total = ts[0].clone()
        for t in ts[1:]:
            total += t
        for t in ts:
            t.copy_(total)

    def _plan(self):
        flags = [b.halo_flags() for b in self.backends]
        self.sync()
        self._sum(flags)
        self.sync()
        for b in self.backends:
            b.halo_plan_async()
        self.planned = True

    def begin(self):
        if not self.planned:
            self._plan()
        if self.peer:                             # every send is enqueued before any recv starts to wait
            return [b.halo_peer_send() for b in self.backends][0]
        self._packed = [b.halo_pack() for b in self.backends]
        return 0 if self._packed[0] is None else self._packed[0].numel() // 256

    def finish(self):
        if self.peer:
            self.sync()                           # one GPU, one host thread: no recv may spin ahead of a send
            for b in self.backends:
                b.halo_peer_recv()
            self._plan()
            return
        self.sync()
        if self._packed[0] is not None:
            self._sum(self._packed)
        self.sync()
        for b in self.backends:
            b.halo_unpack()
        self._plan()

    def exchange(self):
        n = self.begin()
        self.finish()
        return n


def substep(solver, state, exchange, dt, mpm_bc=-1, sort=False, overlap=False):
    """One no-contact substep of the driver loop (deformable_driver.h:240-260 / cuda_mpm_test.cc:66-72)
    on this rank's part of the scene.  With `overlap`, UpdateGrid and GridToParticle run in two parts
    around the exchange: everything that does not touch an exchanged block while the all-reduce is in
    flight, the rest after it (same result as the plain order).  Measured on 2 B200s (1M particles per
    rank): the all-reduce (~35 us) disappears from the critical path, but running G2P as two launches
    costs ~30 us (a chunk of 128 particles is ~25 us of work for its warp, so the second, smaller launch
    is all tail), so the plain order is the default; the split pays once the collective is slower than that."""
    solver.RebuildMapping(state, sort)
    solver.CalcFemStateAndForce(state, dt)
    solver.ParticleToGrid(state, dt)
    if exchange is None:
        solver.UpdateGrid(state, mpm_bc)
        solver.GridToParticle(state, dt)
    elif overlap:
        exchange.begin()
        state.update_grid_part(mpm_bc, 0)
        state.g2p_part(dt, 0)
        exchange.finish()
        state.update_grid_part(mpm_bc, 1)
        state.g2p_part(dt, 1)
    else:
        exchange.exchange()
        solver.UpdateGrid(state, mpm_bc)
        solver.GridToParticle(state, dt)


# ---------------------------------------------------------------------------------------------------------------
# Position-owned slabs with particle migration.  Volumetric particles only: cloth
# meshes stay whole (partition_meshes above).
# ---------------------------------------------------------------------------------------------------------------
def slab_edges(positions, world, axis=0):
    """Interior edges of `world` slabs along `axis` holding equal particle counts (the initial partition):
    world - 1 ascending coordinates; rank r owns [edge[r-1], edge[r])."""
    x = np.sort(np.asarray(positions, np.float64).reshape(-1, 3)[:, axis])
    return np.array([x[(len(x) * r) // world] for r in range(1, world)], np.float64)


def park_points(park, n, start=0):
    """n scattered points of the parking box (the sequence migrate_out uses, csrc/migrate.cuh park_position)."""
    park = np.asarray(park, np.float64).reshape(6)
    a = np.array([0.8191725133961645, 0.6710436067037893, 0.5497004779019703])
    t = (np.arange(start, start + n, dtype=np.float64)[:, None] + 1.0) * a
    return park[:3] + (t - np.floor(t)) * (park[3:] - park[:3])


def owner_of(x, edges):
    return np.searchsorted(np.asarray(edges, np.float64), np.asarray(x, np.float64), side="right")


class MigrationOverflow(RuntimeError):
    """Raised by every rank alike, before any particle was taken out of a state, when some slab has fewer spare
    slots than particles heading for it."""


class SlabMigrator:
    """Keeps the volumetric particles of ONE rank inside its slab [edges[rank-1] - margin, edges[rank] + margin).

    The state must have been created with spare inactive particles (AddParticles with volume 0, scattered over the
    parking box `park` = lo xyz, hi xyz inside the rank's own slab: see park_points).  `gid` maps
    the state's local particle ids to scene-wide ids (-1: inactive) so that results can be compared across
    decompositions.  migrate() is collective: every rank calls it between GridToParticle and the next RebuildMapping,
    every `every` substeps; `all_gather` is torch.distributed.all_gather_object-like (list of per-rank objects).
    Ownership never changes a result (shared grid blocks are summed by the exchange); the margin only keeps particles
    that oscillate around an edge from bouncing between ranks."""

    def __init__(self, state, rank, world, edges, gid, park, axis=0, margin=0.0, cap=1 << 16, all_gather=None):
        self.state, self.rank, self.world = state, rank, world
        self.edges = np.asarray(edges, np.float64)
        self.gid = np.asarray(gid, np.int64).copy()
        self.park = np.asarray(park, np.float64).reshape(6)   # parking box of the inactive particles: lo xyz, hi xyz
        self.axis, self.margin, self.cap = axis, float(margin), int(cap)
        self.all_gather = all_gather
        self.moved_out = self.moved_in = 0

    def bounds(self):
        lo = -np.inf if self.rank == 0 else self.edges[self.rank - 1] - self.margin
        hi = np.inf if self.rank == self.world - 1 else self.edges[self.rank] + self.margin
        big = 1e30
        return max(lo, -big), min(hi, big)

    def extract(self):
        """-> {destination rank: (records, gids)} of the particles that left this rank's slab."""
        lo, hi = self.bounds()
        rec, pids = self.state.migrate_out(self.axis, lo, hi, self.park, self.cap)
        gids = self.gid[pids]
        self.gid[pids] = -1
        dest = owner_of(rec[:, self.axis], self.edges) if len(pids) else np.zeros(0, np.int64)
        self.moved_out += len(pids)
        return {int(d): (rec[dest == d], gids[dest == d]) for d in np.unique(dest)}

    def insert(self, parcels):
        """parcels: list over source ranks of {destination: (records, gids)}."""
        recs = [p[self.rank][0] for p in parcels if self.rank in p]
        gids = [p[self.rank][1] for p in parcels if self.rank in p]
        if not recs:
            return 0
        rec, g = np.concatenate(recs), np.concatenate(gids)
        order = np.argsort(g, kind="stable")          # arrival order must not depend on who sent first
        pids = self.state.migrate_in(rec[order])
        self.gid[pids] = g[order]
        self.moved_in += len(pids)
        return len(pids)

    def check_room(self):
        """Collective dry run: every rank counts its leavers per destination and its free slots (migrate_count, nothing
        is changed), the counts are gathered, and EVERY rank raises MigrationOverflow -- from the same numbers, so
        together -- if some destination cannot seat its arrivals.  Without this a short destination would fail alone,
        after the sources had already parked the leavers, and the other ranks would hang in the next collective.
        Also grows this rank's record buffer so that one migrate_out call takes all its leavers."""
        lo, hi = self.bounds()
        leavers, free = self.state.migrate_count(self.axis, lo, hi, self.edges)
        row = np.concatenate([[self.rank, free], leavers]).astype(np.float64)[None, :]
        if getattr(self.all_gather, "packed", False):
            if row.shape[1] > 28:
                raise ValueError("the packed transport carries at most 26 slabs")
            table = self.all_gather(np.pad(row, ((0, 0), (0, 28 - row.shape[1]))))[:, :row.shape[1]]
        else:
            table = np.concatenate(self.all_gather(row))
        table = table[np.argsort(table[:, 0])]
        free_of, out_of = table[:, 1], table[:, 2:].sum(axis=1)
        arrivals = table[:, 2:].sum(axis=0)
        # a destination's own leavers free their slots before the arrivals are seated (extract runs before insert)
        short = np.nonzero(arrivals > free_of + out_of)[0]
        if len(short):
            d = int(short[0])
            raise MigrationOverflow("slab %d would receive %d particles but has %d free slots (+%d of its own leaving): "
                                    "nothing was moved; create the states with more spare particles, migrate more often "
                                    "or move the slab edges" % (d, arrivals[d], free_of[d], out_of[d]))
        self.cap = max(self.cap, int(leavers.sum()))
        return table

    def migrate(self):
        import time
        if self.all_gather is not None:
            self.check_room()
        t0 = time.perf_counter()
        out = self.extract()
        t1 = time.perf_counter()
        self.t_extract = getattr(self, "t_extract", 0.0) + (t1 - t0)
        if getattr(self.all_gather, "packed", False):
            # NCCL path: the packed records travel as ONE padded tensor all-gather (rows = destination, scene-wide id,
            # 26 scalars) after a MAX all-reduce of the counts -- two small collectives instead of pickled objects
            rows = [np.concatenate([np.full((len(g), 1), d, np.float64), g[:, None].astype(np.float64), r.astype(np.float64)], axis=1)
                    for d, (r, g) in out.items()]
            mine = np.concatenate(rows) if rows else np.zeros((0, 28))
            everything = self.all_gather(mine)
            t2 = time.perf_counter()
            sel = everything[everything[:, 0] == self.rank]
            moved = len(everything)
            if len(sel):
                order = np.argsort(sel[:, 1], kind="stable")
                pids = self.state.migrate_in(sel[order, 2:])
                self.gid[pids] = sel[order, 1].astype(np.int64)
                self.moved_in += len(pids)
            t3 = time.perf_counter()
            self.t_comm = getattr(self, "t_comm", 0.0) + (t2 - t1)
            self.t_insert = getattr(self, "t_insert", 0.0) + (t3 - t2)
            return moved
        parcels = self.all_gather(out)
        self.insert(parcels)
        return sum(len(p[d][1]) for p in parcels for d in p)   # particles that changed hands, scene-wide


def dist_all_gather(group=None):
    import torch.distributed as dist

    def gather(obj):
        out = [None] * dist.get_world_size(group)
        dist.all_gather_object(out, obj, group=group)
        return out
    return gather


def dist_all_gather_packed(device, group=None):
    """all_gather for SlabMigrator over NCCL tensors: f(rows (n, 28) float64) -> the rows of all ranks, concatenated."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)

    def gather(rows):
        n = torch.tensor([rows.shape[0]], dtype=torch.int64, device=device)
        dist.all_reduce(n, op=dist.ReduceOp.MAX, group=group)
        m = int(n.item())
        if m == 0:
            return np.zeros((0, rows.shape[1]))
        buf = torch.full((m, rows.shape[1]), -1.0, dtype=torch.float64, device=device)
        if rows.shape[0]:
            buf[:rows.shape[0]] = torch.from_numpy(np.ascontiguousarray(rows)).to(device)
        out = torch.empty((world, m, rows.shape[1]), dtype=torch.float64, device=device)
        dist.all_gather_into_tensor(out, buf, group=group)
        allr = out.reshape(-1, rows.shape[1]).cpu().numpy()
        return allr[allr[:, 0] >= 0]
    gather.packed = True
    return gather


def migrate_in_process(migrators, exchange=None):
    """The same for several states living in one process (tests).  Returns the number of particles moved."""
    rows = []
    for m in migrators:     # the dry run of SlabMigrator.check_room, all ranks at hand
        lo, hi = m.bounds()
        leavers, free = m.state.migrate_count(m.axis, lo, hi, m.edges)
        rows.append((free, leavers))
        m.cap = max(m.cap, int(leavers.sum()))
    arrivals = np.sum([lv for _, lv in rows], axis=0)
    for d, (free, leavers) in enumerate(rows):
        if arrivals[d] > free + leavers.sum():
            raise MigrationOverflow("slab %d would receive %d particles but has %d free slots: nothing was moved"
                                    % (d, arrivals[d], free))
    parcels = [m.extract() for m in migrators]
    for m in migrators:
        m.insert(parcels)
    moved = sum(len(p[d][1]) for p in parcels for d in p)
    if moved and exchange is not None:
        exchange.planned = False          # the shared-block plan was made for the old ownership: plan afresh
        for b in getattr(exchange, "backends", []):
            if hasattr(b, "reset_plan"):
                b.reset_plan()
    return moved
