"""Test-only backend that runs the shared-block exchange protocol (drake_b200/slab.py) on the CPU oracle,
so that the host side of the multi-GPU path (partitioning, the order and content of the two
all-reduces, the deterministic shared-block list) is exercised with gloo and without a GPU."""
import numpy as np
import torch


class OracleHaloBackend:
    def __init__(self, orc):
        self.orc = orc
        self.own = None
        self.shared = None
        self._flags = None
        self._packed = None

    def halo_flags(self):
        self.own = self.orc.view("grid_touched_flags") != 0
        self._flags = torch.from_numpy(self.own.astype(np.uint8))
        return self._flags

    def halo_plan(self):
        self.shared = np.nonzero(self._flags.numpy() > 1)[0]          # ascending block ids
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
