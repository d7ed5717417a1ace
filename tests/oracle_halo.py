"""Test-only backend that runs the shared-block exchange protocol (drake_b200/slab.py) on the CPU oracle,
so that the host side of the multi-GPU path (partitioning, the order and content of the two
all-reduces, the deterministic shared-block list) is exercised with gloo and without a GPU."""
import numpy as np
import torch


def _contract(v):
    v = v & 0x09249249
    v = (v ^ (v >> 2)) & 0x030C30C3
    v = (v ^ (v >> 4)) & 0x0300F00F
    v = (v ^ (v >> 8)) & 0xFF0000FF
    v = (v ^ (v >> 16)) & 0x000003FF
    return v


def dilate_blocks(flags):
    """flags[n_blocks] (Morton block ids, x most significant) -> flags of the 3x3x3 block neighbourhood."""
    n = flags.size
    side = round(n ** (1.0 / 3.0))
    ids = np.arange(n, dtype=np.uint32)
    bx, by, bz = _contract(ids >> 2), _contract(ids >> 1), _contract(ids)
    dense = np.zeros((side + 2,) * 3, bool)
    dense[bx + 1, by + 1, bz + 1] = flags
    out = np.zeros_like(dense)
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                out |= np.roll(dense, (dx, dy, dz), axis=(0, 1, 2))
    return out[bx + 1, by + 1, bz + 1]


def reachable_blocks(mass):
    """The library's planning rule (csrc/halo.cuh): blocks holding a cell within one CELL of a cell with
    mass.  mass[n_cells] in block-major order -> bool[n_blocks]."""
    n_blocks = mass.size // 64
    side = round(n_blocks ** (1.0 / 3.0))
    ids = np.arange(n_blocks, dtype=np.uint32)
    bx, by, bz = _contract(ids >> 2), _contract(ids >> 1), _contract(ids)
    occ = (mass.reshape(n_blocks, 4, 4, 4) > 0)
    dense = np.zeros((side * 4 + 2,) * 3, bool)
    for lx in range(4):
        for ly in range(4):
            for lz in range(4):
                dense[bx * 4 + lx + 1, by * 4 + ly + 1, bz * 4 + lz + 1] = occ[:, lx, ly, lz]
    out = np.zeros_like(dense)
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                out |= np.roll(dense, (dx, dy, dz), axis=(0, 1, 2))
    out = out[1:-1, 1:-1, 1:-1].reshape(side, 4, side, 4, side, 4).any(axis=(1, 3, 5))
    return out[bx, by, bz]


_WINDOWS = {}   # "device pointer" -> window of an oracle backend in peer mode (same-process attach)


class _Window:
    """CPU stand-in for a peer-memory receive window (csrc/halo_window.cuh): flag[world], present and data
    [2][world][cap]."""

    def __init__(self, world, cap, dtype):
        self.flag = np.zeros(world, np.uint32)
        self.present = np.zeros((2, world, cap), np.uint32)
        self.data = np.zeros((2, world, cap, 4, 64), dtype)


class OracleHaloBackend:
    """Also mirrors the peer-mode calls of GpuMpmState (halo_peer_init / window / attach / send / recv), with
    numpy windows shared by reference, so that slab.InProcessExchange(peer=True) runs on the CPU."""

    def __init__(self, orc):
        self.orc = orc
        self.state = self           # slab.py reaches the peer set-up calls through backend.state
        self.peer = False
        self.planned_own = None     # own dilated flags of the previous plan (containment check)
        self.next_shared = None
        self.shared = None
        self.have_mass = False
        self.own_mass = None
        self._flags = None
        self._packed = None

    def _touched(self):
        return self.orc.view("grid_touched_flags") != 0

    def reset_plan(self):
        """After particles changed hands (slab.SlabMigrator): forget the plan made for the old ownership."""
        self.planned_own = None
        self.next_shared = None
        self.have_mass = False

    def halo_flags(self):
        touched = self._touched()
        if self.planned_own is not None:
            assert not np.any(touched & ~self.planned_own), "touched block outside the planned dilated set"
        if self.have_mass:      # after the first halo_pack: the tight rule, from this rank's own grid masses
            self.planned_own = touched | reachable_blocks(self.own_mass)
        else:                   # start-up (no assembled masses yet): full block dilation
            self.planned_own = dilate_blocks(touched)
        contribution = (1 << self.rank) if self.peer else 1
        self._flags = torch.from_numpy(self.planned_own.astype(np.uint8) * np.uint8(contribution))
        return self._flags

    def halo_plan_async(self):
        f = self._flags.numpy()
        if self.peer:               # flags = sharer set: shared = more than one bit
            self.next_shared = np.nonzero((f & (f - 1)) != 0)[0]
            self.next_sharers = f[self.next_shared].copy()
        else:
            self.next_shared = np.nonzero(f > 1)[0]                   # ascending block ids

    # ---- peer mode (csrc/halo.cuh, second half) ------------------------------------------------------
    def halo_peer_init(self, rank, world, cap_blocks=8192):
        assert not self.peer and 1 <= world <= 8 and rank < world
        self.peer, self.rank, self.world, self.cap, self.xchg = True, rank, world, cap_blocks, 0
        self.windows = [None] * world
        self.windows[rank] = _Window(world, cap_blocks, self.orc.dtype)
        _WINDOWS[id(self.windows[rank])] = self.windows[rank]

    def halo_peer_window(self):
        return b"\0" * 64, id(self.windows[self.rank])

    def halo_peer_attach(self, peer, ipc_handle=None, pointer=None):
        assert self.peer and peer != self.rank and self.windows[peer] is None and pointer is not None
        self.windows[peer] = _WINDOWS[pointer]

    def halo_peer_send(self):
        assert all(w is not None for w in self.windows), "not every peer window is attached"
        self.shared, self.sharers = self.next_shared, self.next_sharers
        assert self.shared.size <= self.cap
        self.own = self._touched()
        self.own_mass = self.orc.view("grid_masses").copy()
        self.have_mass = True
        self.xchg += 1
        par = self.xchg & 1
        mass = self.orc.view("grid_masses").reshape(-1, 64)
        mom = self.orc.view("grid_momentum").reshape(-1, 64, 3)
        peers = 0
        for i, g in enumerate(self.shared):
            who = int(self.sharers[i])
            if who >> self.rank & 1:
                peers |= who
            if not self.own[g]:
                continue                                  # an untouched block sends nothing
            for q in range(self.world):
                if q != self.rank and who >> q & 1:
                    w = self.windows[q]
                    w.data[par, self.rank, i, 0] = mass[g]
                    w.data[par, self.rank, i, 1:] = mom[g].T
                    w.present[par, self.rank, i] = self.xchg
        for q in range(self.world):
            if q != self.rank and peers >> q & 1:
                self.windows[q].flag[self.rank] = self.xchg
        return int(self.shared.size)

    def halo_peer_recv(self):
        par, w = self.xchg & 1, self.windows[self.rank]
        mass = self.orc.view("grid_masses").reshape(-1, 64)
        mom = self.orc.view("grid_momentum").reshape(-1, 64, 3)
        for i, g in enumerate(self.shared):
            if not self.own[g]:
                continue
            who = int(self.sharers[i])
            acc = None
            for q in range(self.world):                   # ascending rank order: the same bits on every sharer
                if not who >> q & 1:
                    continue
                if q == self.rank:
                    v = np.concatenate([mass[g][None, :], mom[g].T])
                else:
                    assert w.flag[q] >= self.xchg, "receive before the sharer delivered"
                    v = w.data[par, q, i] if w.present[par, q, i] == self.xchg else np.zeros((4, 64), self.orc.dtype)
                acc = v.copy() if acc is None else acc + v
            mass[g] = acc[0]
            mom[g] = acc[1:].T

    def halo_pack(self):
        self.shared = self.next_shared
        self.own = self._touched()
        self.own_mass = self.orc.view("grid_masses").copy()   # before the sums come back
        self.have_mass = True
        if self.shared.size == 0:
            self._packed = None
            return None
        mass = self.orc.view("grid_masses").reshape(-1, 64)
        mom = self.orc.view("grid_momentum").reshape(-1, 64, 3)
        packed = np.zeros((self.shared.size, 4, 64), self.orc.dtype)
        mine = self.own[self.shared]
        packed[mine, 0, :] = mass[self.shared[mine]]
        packed[mine, 1:, :] = mom[self.shared[mine]].transpose(0, 2, 1)
        self._packed = torch.from_numpy(packed.reshape(-1))
        return self._packed

    def halo_unpack(self):
        if self._packed is None:
            return
        packed = self._packed.numpy().reshape(-1, 4, 64)
        mine = self.own[self.shared]
        mass = self.orc.view("grid_masses").reshape(-1, 64)
        mom = self.orc.view("grid_momentum").reshape(-1, 64, 3)
        mass[self.shared[mine]] = packed[mine, 0, :]
        mom[self.shared[mine]] = packed[mine, 1:, :].transpose(0, 2, 1)


def oracle_substep(orc, exchange, dt, mpm_bc=-1, sort=False):
    orc.RebuildMapping(sort)
    orc.CalcFemStateAndForce(dt)
    orc.ParticleToGrid(dt)
    if exchange is not None:
        exchange()
    orc.UpdateGrid(mpm_bc)
    orc.GridToParticle(dt)
