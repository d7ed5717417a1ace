"""Host-side mirror of the reference's multibody/gpu_mpm interface on top of the C ABI.

Same names, argument meaning and call order as the reference's C++ classes so that tests and
drivers read like the reference's own (cuda_mpm_test.cc:18-79, deformable_driver.h:221-271):

  GpuMpmState   <- GpuMpmState<T>   (cuda_mpm_model.cuh:37-260)
  GpuMpmSolver  <- GpuMpmSolver<T>  (cuda_mpm_solver.cuh:19-35)
  MpmParticleContactPairs, MpmConfigParams, CpuMpmModel, MpmPortData, ExternalSpatialForce
                <- cpu_mpm_model.h:16-114, cuda_mpm_model.cuh:20-35

The C++ twin of this file (header-only, for linking into Drake itself) is
include/multibody/gpu_mpm/.  All compute happens in libgmpm_b200.so on the GPU; there is no CPU path.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import ContactStats, GmpmError, Params, check, default_params  # noqa: F401


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class MpmConfigParams:
    """cpu_mpm_model.h:17-26."""

    def __init__(self, substep_dt=1e-3, write_files=False, contact_stiffness=1e5, contact_damping=0.0,
                 contact_friction_mu=0.0, contact_query_frequency=1, mpm_bc=-1, exact_line_search=False):
        self.substep_dt = substep_dt
        self.write_files = write_files
        self.contact_stiffness = contact_stiffness
        self.contact_damping = contact_damping
        self.contact_friction_mu = contact_friction_mu
        self.contact_query_frequency = contact_query_frequency
        self.mpm_bc = mpm_bc
        self.exact_line_search = exact_line_search


class CpuMpmModel:
    """cpu_mpm_model.h:33-40: host-side holder of the initial cloth and the config."""

    def __init__(self):
        self.cloth_pos = []
        self.cloth_vel = []
        self.cloth_indices = []
        self.config = MpmConfigParams()


class MpmPortData:
    """cpu_mpm_model.h:46-49."""

    def __init__(self, pos=None, indices=None):
        self.pos = pos
        self.indices = indices


class ExternalSpatialForce:
    """cuda_mpm_model.cuh:20-35."""

    def __init__(self, dtype):
        self._dtype = dtype
        self.resize(0)

    def size(self):
        return self.p_BoBq_B.shape[0]

    def resize(self, n):
        self.p_BoBq_B = np.zeros((n, 3), self._dtype)
        self.F_Bq_W_tau = np.zeros((n, 3), self._dtype)
        self.F_Bq_W_f = np.zeros((n, 3), self._dtype)


class MpmParticleContactPairs:
    """cpu_mpm_model.h:74-114 (SoA of the particles that fell inside a rigid geometry)."""

    def __init__(self, dtype=np.float64):
        self._dtype = np.dtype(dtype)
        self.clear()

    def clear(self):
        self.particle_in_contact_index = []
        self.non_mpm_id = []
        self.penetration_distance = []
        self.normal = []
        self.particle_in_contact_position = []
        self.rigid_v = []
        self.rigid_p_WB = []

    def push_back(self, particle_index, non_mpm_id, penetration_distance, normal, position, rigid_v, rigid_p_WB):
        self.particle_in_contact_index.append(int(particle_index))
        self.non_mpm_id.append(int(non_mpm_id))
        self.penetration_distance.append(float(penetration_distance))
        self.normal.append(np.asarray(normal, self._dtype))
        self.particle_in_contact_position.append(np.asarray(position, self._dtype))
        self.rigid_v.append(np.asarray(rigid_v, self._dtype))
        self.rigid_p_WB.append(np.asarray(rigid_p_WB, self._dtype))

    def extend(self, particle_index, non_mpm_id, penetration_distance, normal, position, rigid_v, rigid_p_WB):
        """Vectorised push_back (arrays of equal length)."""
        self.particle_in_contact_index = list(self.particle_in_contact_index) + list(np.asarray(particle_index))
        self.non_mpm_id = list(self.non_mpm_id) + list(np.asarray(non_mpm_id))
        self.penetration_distance = list(self.penetration_distance) + list(np.asarray(penetration_distance))
        for dst, src in ((self.normal, normal), (self.particle_in_contact_position, position),
                         (self.rigid_v, rigid_v), (self.rigid_p_WB, rigid_p_WB)):
            dst.extend(np.asarray(src, self._dtype).reshape(-1, 3))

    def size(self):
        return len(self.non_mpm_id)

    def arrays(self):
        n = self.size()
        f = lambda x, w: np.ascontiguousarray(np.asarray(x, self._dtype).reshape(n, w) if n else np.zeros((0, w), self._dtype))
        return (np.ascontiguousarray(self.particle_in_contact_index, np.uint32),
                np.ascontiguousarray(self.non_mpm_id, np.uint32), f(self.penetration_distance, 1).reshape(n),
                f(self.normal, 3), f(self.particle_in_contact_position, 3), f(self.rigid_v, 3), f(self.rigid_p_WB, 3))


# buffer name -> (kind, extent, components); kind T = scalar, i = int32, u = uint32
_BUFFERS = {
    "positions": ("T", "N", 3), "velocities": ("T", "N", 3), "volumes": ("T", "N", 1),
    "affine_matrices": ("T", "N", 9), "pids": ("i", "N", 1), "sort_keys": ("u", "N", 1), "sort_ids": ("u", "N", 1),
    "forces": ("T", "N", 3), "taus": ("T", "N", 9), "index_mappings": ("i", "N", 1), "indices": ("i", "F", 3),
    "Dm_inverses": ("T", "F", 4), "deformation_gradients": ("T", "F", 9), "grid_masses": ("T", "cells", 1),
    "grid_momentum": ("T", "cells", 3), "grid_v_star": ("T", "cells", 3), "grid_touched_flags": ("u", "blocks", 1),
    "grid_touched_ids": ("u", "blocks", 1), "grid_Hess": ("T", "cells", 9), "grid_Grad": ("T", "cells", 3),
    "grid_Dir": ("T", "cells", 3), "contact_vel": ("T", "Nc", 3), "contact_vel0": ("T", "Nc", 3),
    "contact_sort_keys": ("u", "Nc", 1), "contact_mpm_id": ("u", "Nc", 1), "contact_rigid_id": ("u", "Nc", 1),
    "contact_dist": ("T", "Nc", 1), "contact_normal": ("T", "Nc", 3), "contact_pos": ("T", "Nc", 3),
    "contact_rigid_v": ("T", "Nc", 3), "contact_rigid_p": ("T", "Nc", 3),
    "particle_F": ("T", "Nv", 9),   # deformation gradients of the volumetric particles, by particle id (extension)
}


class GpuMpmState:
    """GpuMpmState<T> (cuda_mpm_model.cuh:37-260).  dtype float32 = the reference's product build
    (config::GpuT, settings.h:37); float64 = the same algorithms at double."""

    def __init__(self, dtype=np.float64, params=None, **param_overrides):
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
            raise TypeError("dtype must be float32 or float64")
        self._suf = "f32" if self.dtype == np.float32 else "f64"
        self._ct = C.c_float if self._suf == "f32" else C.c_double
        self._L = _lib.lib()
        self.params = params if params is not None else default_params(**param_overrides)
        h = C.c_void_p()
        check(self._f("create")(C.byref(self.params), C.byref(h)))
        self._h = h
        self._positions_host = np.zeros((0, 3), self.dtype)
        self._positions_pending = False
        self._external_forces_host = ExternalSpatialForce(self.dtype)

    def _f(self, name):
        return getattr(self._L, "gmpm_%s_%s" % (name, self._suf))

    # --- sizes
    def n_verts(self):
        return self._f("n_verts")(self._h)

    def n_faces(self):
        return self._f("n_faces")(self._h)

    def n_particles(self):
        return self._f("n_particles")(self._h)

    def num_contacts(self):
        return self._f("num_contacts")(self._h)

    def n_cells(self):
        return self._f("n_cells")(self._h)

    def grid_touched_cnt_host(self):
        c = C.c_uint32()
        check(self._f("grid_touched_cnt_host")(self._h, C.byref(c)))
        return c.value

    def block_counts(self):
        a, b = C.c_uint32(), C.c_uint32()
        check(self._f("block_counts")(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def device_bytes(self):
        return self._f("device_bytes")(self._h)

    # --- checkpoint / restore (gmpm_c.h; the reference has none)
    def save_state(self):
        """-> bytes: positions, velocities, volumes, affine matrices, pids, F, Dm^-1 (reference layout)."""
        n = self._f("checkpoint_bytes")(self._h)
        buf = np.zeros(n, np.uint8)
        w = C.c_size_t()
        check(self._f("save_state")(self._h, _ptr(buf), n, C.byref(w)))
        return buf[:w.value].tobytes()

    def load_state(self, blob):
        buf = np.frombuffer(blob, np.uint8)
        check(self._f("load_state")(self._h, C.c_void_p(buf.ctypes.data), buf.size))
        self._positions_pending = False

    # --- multi-GPU shared-block exchange (gmpm_c.h, "multi-GPU extension"); driven by drake_b200.slab
    def stream_handle(self):
        p = C.c_void_p()
        check(self._f("stream")(self._h, C.byref(p)))
        return p.value or 0

    def halo_reset_plan(self):
        check(self._f("halo_reset_plan")(self._h))

    def halo_flags(self):
        """-> (device pointer of uint8[n_blocks], n_blocks)"""
        p, n = C.c_void_p(), C.c_size_t()
        check(self._f("halo_flags")(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def halo_stream_handle(self):
        p = C.c_void_p()
        check(self._f("halo_stream")(self._h, C.byref(p)))
        return p.value or 0

    def halo_plan_async(self):
        check(self._f("halo_plan_async")(self._h))

    def halo_pack(self):
        """-> (n_shared, device pointer of T[n_shared][4][64])"""
        n, p = C.c_uint32(), C.c_void_p()
        check(self._f("halo_pack")(self._h, C.byref(n), C.byref(p)))
        return n.value, p.value

    def halo_unpack(self):
        check(self._f("halo_unpack")(self._h))

    # peer mode: the exchange through receive windows in peer memory (gmpm_c.h)
    def halo_peer_init(self, rank, world, cap_blocks=8192):
        check(self._f("halo_peer_init")(self._h, rank, world, cap_blocks))

    def halo_peer_window(self):
        """-> (64-byte CUDA IPC handle, raw device pointer) of this rank's receive window."""
        h = C.create_string_buffer(64)
        p = C.c_void_p()
        check(self._f("halo_peer_window")(self._h, h, C.byref(p)))
        return h.raw, p.value

    def halo_peer_attach(self, peer, ipc_handle=None, pointer=None):
        check(self._f("halo_peer_attach")(self._h, peer, ipc_handle, pointer))

    def halo_peer_send(self):
        n = C.c_uint32()
        check(self._f("halo_peer_send")(self._h, C.byref(n)))
        return n.value

    def halo_peer_recv(self):
        check(self._f("halo_peer_recv")(self._h))

    def update_grid_part(self, mpm_bc, part):
        check(self._f("update_grid_part")(self._h, int(mpm_bc), int(part)))

    def g2p_part(self, dt, part):
        check(self._f("g2p_part")(self._h, self._ct(dt), int(part)))

    # --- host mirrors
    def positions_host(self):
        """cuda_mpm_model.cuh:116-117.  Waits for the snapshot SyncParticleStateToCpu started and returns
        a read-only (N, 3) view of the library's pinned mirror (or the caller's array of a blocking sync)."""
        if self._positions_pending:
            p = C.c_void_p()
            check(self._f("positions_host")(self._h, C.byref(p)))
            n = self.n_particles()
            ctype = C.c_double if self.dtype == np.float64 else C.c_float
            arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(ctype)), shape=(n, 3)) if p.value and n else np.zeros((0, 3), self.dtype)
            arr.flags.writeable = False
            self._positions_host = arr
            self._positions_pending = False
        return self._positions_host

    def SetContactQueryBounds(self, boxes):
        """Extension: axis-aligned bounds [(lo, hi), ...] of the rigid geometry the host contact query tests against.
        The async snapshot then moves only the particles inside one of them to the host; positions_host() of the
        others reads 1e30.  `None` restores the full copy."""
        if boxes is None:
            check(self._f("clear_query_bounds")(self._h))
            return
        arr = np.ascontiguousarray([list(lo) + list(hi) for lo, hi in boxes], self.dtype).reshape(-1, 6)
        check(self._f("set_query_bounds")(self._h, _ptr(arr) if arr.size else None, arr.shape[0]))

    def snapshot_d2h_bytes(self):
        n = C.c_uint64()
        check(self._f("snapshot_d2h_bytes")(self._h, C.byref(n)))
        return n.value

    def external_forces_host(self):
        return self._external_forces_host

    # --- setup
    def AddQRCloth(self, pos, vel, indices):
        pos = np.ascontiguousarray(pos, self.dtype).reshape(-1, 3)
        vel = np.ascontiguousarray(vel, self.dtype).reshape(-1, 3)
        tri = np.ascontiguousarray(indices, np.int32).reshape(-1, 3)
        if vel.shape != pos.shape:
            raise ValueError("pos and vel must have the same shape")
        check(self._f("add_qr_cloth")(self._h, _ptr(pos), _ptr(vel), pos.shape[0], _ptr(tri), tri.shape[0]))

    def AddParticles(self, pos, vel, volume, model="fixed_corotated", youngs=1e5, poisson=0.3, friction_angle_deg=30.0,
                     cohesion=0.0):
        """Extension: particles with their own deformation gradient and a volumetric material
        ("fixed_corotated", "stvk_hencky", "drucker_prager").  `volume`: rest volume per particle (scalar or array)."""
        pos = np.ascontiguousarray(pos, self.dtype).reshape(-1, 3)
        vel = np.ascontiguousarray(vel, self.dtype).reshape(-1, 3)
        vol = np.ascontiguousarray(np.broadcast_to(np.asarray(volume, self.dtype), (pos.shape[0],)))
        if vel.shape != pos.shape:
            raise ValueError("pos and vel must have the same shape")
        m = _lib.Material(_lib.MATERIAL_MODELS[model], 0, float(youngs), float(poisson), float(friction_angle_deg), float(cohesion))
        check(self._f("add_particles")(self._h, _ptr(pos), _ptr(vel), _ptr(vol), pos.shape[0], C.byref(m)))

    MIGRATE_RECORD = 26   # pos 3, vel 3, C 9, volume, F 9, material index

    def migrate_out(self, axis, lo, hi, park, cap):
        """Extension: active volumetric particles outside [lo, hi) along `axis` leave the state.
        -> (records (n, 26), local particle ids (n,)), sorted by id; at most cap."""
        buf = getattr(self, "_migrate_buf", None)       # reused: zero-filling 26 * cap scalars per call costs milliseconds
        if buf is None or buf[0].shape[0] < cap:
            buf = self._migrate_buf = (np.empty((cap, self.MIGRATE_RECORD), self.dtype), np.empty(cap, np.int32))
        rec, pids = buf
        park = np.ascontiguousarray(park, self.dtype)
        n = C.c_size_t()
        check(self._f("migrate_out")(self._h, int(axis), self._ct(lo), self._ct(hi), _ptr(park), _ptr(rec), _ptr(pids),
                                     cap, C.byref(n)))
        return rec[:n.value].copy(), pids[:n.value].copy()

    def migrate_count(self, axis, lo, hi, edges):
        """Extension: dry run of migrate_out.  -> (leavers per destination slab (len(edges) + 1,), free slots)."""
        e = np.ascontiguousarray(edges, np.float64).reshape(-1)
        out = np.zeros(e.size + 1, np.uint64)
        free = C.c_size_t()
        check(self._f("migrate_count")(self._h, int(axis), self._ct(lo), self._ct(hi), _ptr(e), e.size, _ptr(out),
                                       C.byref(free)))
        return out.astype(np.int64), int(free.value)

    def migrate_in(self, records):
        """Extension: the records take over inactive particles (volume 0).  -> the local particle ids they took."""
        rec = np.ascontiguousarray(records, self.dtype).reshape(-1, self.MIGRATE_RECORD)
        pids = np.zeros(rec.shape[0], np.int32)
        if rec.shape[0]:
            check(self._f("migrate_in")(self._h, _ptr(rec), rec.shape[0], _ptr(pids)))
        return pids

    def n_volumetric(self):
        return self._f("n_volumetric")(self._h)

    def Finalize(self):
        check(self._f("finalize")(self._h))

    def Destroy(self):
        if self._h:
            check(self._f("destroy")(self._h))
            self._h = None

    def __del__(self):
        try:
            self.Destroy()
        except Exception:
            pass

    def DumpCpuState(self):
        pos = np.zeros((self.n_verts(), 3), self.dtype)
        tri = np.zeros((self.n_faces(), 3), np.int32)
        check(self._f("dump_cpu_state")(self._h, _ptr(pos), _ptr(tri)))
        return pos, tri.reshape(-1)

    def ReallocateExternelBodies(self, n):
        check(self._f("reallocate_external_bodies")(self._h, int(n)))
        self._n_bodies = int(n)

    def ExternelBodyForceToHost(self):
        n = getattr(self, "_n_bodies", 0)
        ext = self._external_forces_host
        if ext.size() != n:
            ext.resize(n)
        check(self._f("external_body_force_to_host")(self._h, _ptr(ext.F_Bq_W_tau), _ptr(ext.F_Bq_W_f)))

    # --- raw-accessor replacement (reference layout, visible order)
    def read(self, name):
        kind, extent, comps = _BUFFERS[name]
        n = {"N": self.n_particles(), "F": self.n_faces(), "cells": self.n_cells(), "blocks": self.n_cells() // 64,
             "Nc": self.num_contacts(), "Nv": self.n_volumetric() if extent == "Nv" else 0}[extent]
        dt = {"T": self.dtype, "i": np.dtype(np.int32), "u": np.dtype(np.uint32)}[kind]
        out = np.zeros(n * comps, dt)
        w = C.c_size_t()
        check(self._f("read")(self._h, name.encode(), _ptr(out), out.nbytes, C.byref(w)))
        assert w.value == out.nbytes, (name, w.value, out.nbytes)
        return out

    # --- instrumentation
    def dev_counters(self):
        """The device-side counters (csrc/kernels.cuh DevCounters) as a dict."""
        out = np.zeros(28, np.uint32)
        w = C.c_size_t()
        check(self._f("read")(self._h, b"dev_counters", _ptr(out), out.nbytes, C.byref(w)))
        names = ["n_pblocks", "n_touched", "n_touched_prev", "work_p2g", "work_g2p", "work_grid", "n_big", "n_small",
                 "n_tiles", "win_lo_x", "win_lo_y", "win_lo_z", "win_hi_x", "win_hi_y", "win_hi_z", "win_valid", "win_violation", "win_fallbacks",
                 "sort_passes"]
        d = dict(zip(names, (int(v) for v in out)))
        d["bin_phase_ns"] = [int(v) for v in out[19:23]]   # window, sort, block ranges, work records (cumulative)
        d["sort_pass0_ns"] = [int(v) for v in out[23:26]]  # end of hist / scan / scatter of pass 0 (cumulative)
        return d

    def kernel_launches(self):
        return int(self._f("kernel_launches")(self._h))

    def set_profiling(self, on):
        check(self._f("set_profiling")(self._h, 1 if on else 0))

    def kernel_time(self, prefix):
        ms, n = C.c_double(), C.c_uint64()
        check(self._f("kernel_time")(self._h, prefix.encode(), C.byref(ms), C.byref(n)))
        return ms.value, int(n.value)

    def timer_mark(self, slot):
        check(self._f("timer_mark")(self._h, int(slot)))

    def timer_elapsed_ms(self, a, b):
        ms = C.c_double()
        check(self._f("timer_elapsed_ms")(self._h, int(a), int(b), C.byref(ms)))
        return ms.value

    def profile_report(self):
        need = self._f("profile_report")(self._h, None, 0)
        buf = C.create_string_buffer(max(int(need), 1))
        self._f("profile_report")(self._h, buf, len(buf))
        out = {}
        for line in buf.value.decode().splitlines():
            name, ms, n = line.split()
            out[name] = (float(ms), int(n))
        return out


class GpuMpmSolver:
    """GpuMpmSolver<T> (cuda_mpm_solver.cuh:19-35): stateless launch logic."""

    def RebuildMapping(self, state, sort):
        check(state._f("rebuild_mapping")(state._h, 1 if sort else 0))

    def CalcFemStateAndForce(self, state, dt):
        check(state._f("calc_fem")(state._h, state._ct(dt)))

    def ParticleToGrid(self, state, dt):
        check(state._f("p2g")(state._h, state._ct(dt)))

    def UpdateGrid(self, state, mpm_bc=-1):
        check(state._f("update_grid")(state._h, int(mpm_bc)))

    def GridToParticle(self, state, dt):
        check(state._f("g2p")(state._h, state._ct(dt)))

    def GpuSync(self, state=None):
        # the reference's GpuSync is a cudaDeviceSynchronize; here work lives on the state's stream
        if state is not None:
            check(state._f("sync")(state._h))

    def SyncParticleStateToCpu(self, state, out=None):
        """cuda_mpm_solver.cu:187-191: positions_host() <- current positions (3N, visible order).
        Without `out` the snapshot is only STARTED here (device gather in stream order, D2H on a copy
        stream into the library's pinned mirror) and `state.positions_host()` waits for it -- the driver
        reads it four solver calls later (deformable_driver.h:243 vs :120-194).  With a caller-owned
        `out` array of shape (N, 3) the call blocks and fills it, like the reference."""
        if out is None:
            check(state._f("sync_particle_state_to_cpu_async")(state._h))
            state._positions_pending = True
            return None
        state._positions_pending = False
        state._positions_host = out
        check(state._f("sync_particle_state_to_cpu")(state._h, C.c_void_p(out.ctypes.data)))
        return out

    def Dump(self, state, filename):
        """cuda_mpm_solver.cu:169-183: OBJ of the cloth vertices and triangles."""
        pos, tri = state.DumpCpuState()
        with open(filename, "w") as f:
            for v in pos:
                f.write("v %g %g %g\n" % (v[0], v[1], v[2]))
            for t in tri.reshape(-1, 3):
                f.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))

    def CopyContactPairs(self, state, pairs):
        ids, rid, dist, nrm, pos, rv, rp = pairs.arrays()
        keep = [a.astype(state.dtype, copy=False) for a in (dist, nrm, pos, rv, rp)]
        keep = [np.ascontiguousarray(a) for a in keep]
        check(state._f("copy_contact_pairs")(state._h, ids.size, _ptr(ids), _ptr(rid), *[_ptr(a) for a in keep]))

    def GenerateContactPairs(self, state, geometries):
        """Device-side CalcMpmContactPairs + CopyContactPairs for analytic rigid geometry (boxes, spheres):
        `geometries` = list of dicts {kind: "box"|"sphere", rigid_id, lo, hi | center, radius, velocity, p_WB}.
        Returns the number of pairs; UpdateContact follows as after CopyContactPairs."""
        arr = self._geometry_array(geometries)
        n = C.c_size_t()
        check(state._f("generate_contact_pairs")(state._h, arr, len(geometries), C.byref(n)))
        return n.value

    @staticmethod
    def _geometry_array(geometries):
        arr = (_lib.AnalyticGeometry * max(1, len(geometries)))()
        for a, g in zip(arr, geometries):
            sphere = g.get("kind", "box") == "sphere"
            a.kind = 1 if sphere else 0
            a.rigid_id = int(g["rigid_id"])
            lo = g["center"] if sphere else g["lo"]
            hi = (g["radius"], 0.0, 0.0) if sphere else g["hi"]
            vel = g.get("velocity", (0.0, 0.0, 0.0))
            pwb = g.get("p_WB", lo if sphere else [0.5 * (x + y) for x, y in zip(lo, hi)])
            for d in range(3):
                a.lo[d], a.hi[d], a.velocity[d], a.p_WB[d] = float(lo[d]), float(hi[d]), float(vel[d]), float(pwb[d])
        return arr

    def RunDriverSubsteps(self, state, n, dt, geometries=(), mpm_bc=-1, sort_every=0, contact_query_frequency=1,
                          friction_mu=0.0, stiffness=0.0, damping=0.0, exact_line_search=False, max_iterations=0,
                          freeze_geometry=False):
        """n substeps of the DeformableDriver loop (deformable_driver.h:240-260) in one ABI call, contact pairs
        from analytic geometry generated on the device every `contact_query_frequency` substeps
        (MpmConfigParams::contact_query_frequency).  Geometries move with their velocity.  -> DriverStats."""
        geometries = list(geometries)
        arr = self._geometry_array(geometries)
        # freeze_geometry: the reference driver's behaviour (rigid poses constant over the substeps of a frame)
        check(state._f("set_driver_geometry_frozen")(state._h, 1 if freeze_geometry else 0))
        st = _lib.DriverStats()
        ct = state._ct
        check(state._f("run_driver_substeps")(state._h, int(n), ct(dt), int(mpm_bc), int(sort_every), arr, len(geometries),
                                              int(contact_query_frequency), ct(friction_mu), ct(stiffness), ct(damping),
                                              1 if exact_line_search else 0, int(max_iterations), C.byref(st)))
        state._positions_pending = False
        return st

    def UpdateContact(self, state, frame, substep, dt, friction_mu, stiffness, damping, dump=False,
                      exact_line_search=False, max_iterations=0):
        st = ContactStats()
        ct = state._ct
        check(state._f("update_contact")(state._h, ct(dt), ct(friction_mu), ct(stiffness), ct(damping),
                                         1 if exact_line_search else 0, int(max_iterations), C.byref(st)))
        return st

    # convenience: n substeps of the no-contact driver loop in one ABI call
    def RunSubsteps(self, state, n, dt, mpm_bc=-1, sort_every=0):
        check(state._f("run_substeps")(state._h, int(n), state._ct(dt), int(mpm_bc), int(sort_every)))


def radix_sort_pairs(keys, ids, num_bit):
    """Stable device sort of (key,id) pairs on key bits [0,num_bit) (radix_sort.cuh:202-286 semantics)."""
    keys = np.ascontiguousarray(keys, np.uint32)
    ids = np.ascontiguousarray(ids, np.uint32)
    ko, io = np.zeros_like(keys), np.zeros_like(ids)
    check(_lib.lib().gmpm_radix_sort_pairs(_ptr(keys), _ptr(ids), keys.size, int(num_bit), _ptr(ko), _ptr(io)))
    return ko, io
