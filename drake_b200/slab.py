"""Multi-GPU driver of the MPM hot path: one process and one GpuMpmState per GPU, scenes partitioned by
whole cloth meshes, grid blocks reached by several ranks summed between ParticleToGrid and UpdateGrid.

The reference is single-GPU (multibody/gpu_mpm/settings.h:40: G_DEVICE_COUNT = 1, never read); this is
the extension SURVEY.md 8(e) plans.  The device side lives in libgmpm_b200.so (csrc/halo.cuh, entry
points gmpm_halo_flags / gmpm_halo_plan / gmpm_halo_unpack); this module is the host plumbing: it owns
no arithmetic, it only orders the two collectives between those calls:

    flags  = backend.halo_flags()          # uint8[n_blocks], 1 = this rank's P2G touched the block
    all_reduce(flags, SUM)                 # NCCL over NVLink (gloo in the CPU tests)
    packed = backend.halo_plan()           # T[n_shared * 256]: own mass + momentum of blocks with flags > 1
    all_reduce(packed, SUM)
    backend.halo_unpack()                  # sums written back where this rank touches the block

`backend` is anything with those three methods returning torch tensors that alias its buffers: the GPU
backend below wraps a GpuMpmState (tensors alias device memory, collectives are enqueued on the state's
own CUDA stream, so the whole substep stays stream-ordered with a single 4-byte read-back in
halo_plan); the CPU tests drive the same protocol class with a backend built on the oracle.
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


class _DevArray:
    """Zero-copy view of a raw device pointer for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


class GpuHaloBackend:
    """The three device steps of one GpuMpmState as torch tensors aliasing the library's buffers."""

    def __init__(self, state):
        import torch
        self.state = state
        self.torch = torch
        self.device = torch.device("cuda", state.params.device if state.params.device >= 0 else torch.cuda.current_device())
        self.stream = torch.cuda.ExternalStream(state.stream_handle(), device=self.device)
        self._typestr = "<f8" if state.dtype == np.float64 else "<f4"
        self.n_shared = 0

    def _view(self, ptr, n, typestr):
        return self.torch.as_tensor(_DevArray(ptr, n, typestr), device=self.device)

    def halo_flags(self):
        ptr, n = self.state.halo_flags()
        return self._view(ptr, n, "|u1")

    def halo_plan(self):
        self.n_shared, ptr = self.state.halo_plan()
        if self.n_shared == 0:
            return None
        return self._view(ptr, self.n_shared * 256, self._typestr)

    def halo_unpack(self):
        self.state.halo_unpack()


class SlabExchange:
    """Runs the exchange protocol over a torch.distributed process group (NCCL on GPUs, gloo on CPU)."""

    def __init__(self, backend, group=None):
        import torch.distributed as dist
        self.backend = backend
        self.dist = dist
        self.group = group
        self.stream = getattr(backend, "stream", None)
        self.last_n_shared = 0
        self.bytes_reduced = 0

    def _all_reduce(self, t):
        if self.stream is not None:
            import torch
            with torch.cuda.stream(self.stream):   # ordered after the state's kernels, before its next ones
                self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        else:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        self.bytes_reduced += t.numel() * t.element_size()

    def exchange(self):
        """Call between ParticleToGrid and UpdateGrid, on every rank."""
        flags = self.backend.halo_flags()
        self._all_reduce(flags)
        packed = self.backend.halo_plan()     # every rank derives the same shared-block list
        self.last_n_shared = 0 if packed is None else packed.numel() // 256
        if packed is not None:
            self._all_reduce(packed)
        self.backend.halo_unpack()
        return self.last_n_shared


def exchange_in_process(backends):
    """The same protocol for several states living in ONE process (tests: two states on one GPU, or
    two oracles on the CPU): the all-reduces become explicit sums over the backends' tensors."""
    flags = [b.halo_flags() for b in backends]
    total = flags[0].clone()
    for f in flags[1:]:
        total += f
    for f in flags:
        f.copy_(total)
    packed = [b.halo_plan() for b in backends]
    if packed[0] is not None:
        total = packed[0].clone()
        for p in packed[1:]:
            total += p
        for p in packed:
            p.copy_(total)
    for b in backends:
        b.halo_unpack()
    return 0 if packed[0] is None else packed[0].numel() // 256


def substep(solver, state, exchange, dt, mpm_bc=-1, sort=False):
    """One no-contact substep of the driver loop (deformable_driver.h:240-260 / cuda_mpm_test.cc:66-72)
    on this rank's part of the scene, with the exchange between P2G and the grid update."""
    solver.RebuildMapping(state, sort)
    solver.CalcFemStateAndForce(state, dt)
    solver.ParticleToGrid(state, dt)
    if exchange is not None:
        exchange.exchange()
    solver.UpdateGrid(state, mpm_bc)
    solver.GridToParticle(state, dt)
