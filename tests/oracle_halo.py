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


class OracleHaloBackend:
    def __init__(self, orc):
        self.orc = orc
        self.planned_own = None     # own dilated flags of the previous plan (containment check)
        self.next_shared = None
        self.shared = None
        self._flags = None
        self._packed = None

    def _touched(self):
        return self.orc.view("grid_touched_flags") != 0

    def halo_flags(self):
        touched = self._touched()
        if self.planned_own is not None:
            assert not np.any(touched & ~self.planned_own), "touched block outside the planned dilated set"
        self.planned_own = dilate_blocks(touched)
        self._flags = torch.from_numpy(self.planned_own.astype(np.uint8))
        return self._flags

    def halo_plan_async(self):
        self.next_shared = np.nonzero(self._flags.numpy() > 1)[0]     # ascending block ids

    def halo_pack(self):
        self.shared = self.next_shared
        self.own = self._touched()
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
