"""ctypes binding of the CPU oracle (oracle/libgmpm_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may import
this.  The product package `drake_b200` never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgmpm_oracle.so")


class OracleParams(C.Structure):
    """Mirror of ogmpm_params (oracle/gmpm_oracle.h); defaults = reference settings.h:37-126."""

    _fields_ = [
        ("domain_bits", C.c_int),
        ("boundary_cells", C.c_int),
        ("gravity_axis", C.c_int),
        ("reserved", C.c_int),
        ("youngs", C.c_double),
        ("poisson", C.c_double),
        ("density", C.c_double),
        ("gamma", C.c_double),
        ("k_normal", C.c_double),
        ("v_blend", C.c_double),
        ("c_f", C.c_double),
        ("sdf_friction", C.c_double),
        ("gravity", C.c_double),
        ("epsv", C.c_double),
        ("dx", C.c_double),
    ]


class ContactStats(C.Structure):
    _fields_ = [
        ("iterations", C.c_int),
        ("grid_dofs", C.c_int),
        ("line_search_evals", C.c_int),
        ("residual", C.c_double),
        ("energy", C.c_double),
    ]


def build(force=False):
    """Compile the C restatement with gcc (seconds).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("gmpm_oracle.c", "gmpm_oracle_impl.inc", "gmpm_oracle_vol.inc", "gmpm_oracle.h")]
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


FAST_FLAGS = ["-std=c11", "-O3", "-march=native", "-fopenmp", "-fno-math-errno"]


def build_fast():
    """Timing-only build of the SAME sources (`-O3 -march=native`, FMA contraction allowed) for bench.py's
    cpu_baseline / --impl reference.  The bit-exact `-O2 -ffp-contract=off` library stays the oracle
    of the parity tests.  `-march=native` binds the file to the host CPU, so it is compiled where it runs (the GPU
    box has gcc) and named after that CPU's feature flags."""
    import hashlib
    try:
        flags = next(l for l in open("/proc/cpuinfo") if l.startswith("flags"))
    except Exception:
        flags = "unknown"
    tag = hashlib.sha1(flags.encode()).hexdigest()[:10]
    out_dir = os.path.join(_HERE, "_fast")
    path = os.path.join(out_dir, "libgmpm_oracle_fast_%s.so" % tag)
    srcs = [os.path.join(_HERE, f) for f in ("gmpm_oracle.c", "gmpm_oracle_impl.inc", "gmpm_oracle_vol.inc", "gmpm_oracle.h")]
    if os.path.exists(path) and os.path.getmtime(path) >= max(os.path.getmtime(s) for s in srcs):
        return path
    os.makedirs(out_dir, exist_ok=True)
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"   # the image's CC wrapper lacks libgomp.spec
    subprocess.check_call([cc, *FAST_FLAGS, "-shared", "-fPIC", srcs[0], "-o", path, "-lm"])
    return path


_lib = None
_lib_fast = None


def lib(fast=False):
    global _lib, _lib_fast
    if fast:
        if _lib_fast is None:
            _lib_fast = C.CDLL(build_fast())
            _declare(_lib_fast)
        return _lib_fast
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _declare(_lib)
    return _lib


_U32P = C.POINTER(C.c_uint32)


def _declare(L):
    L.ogmpm_default_params.argtypes = [C.POINTER(OracleParams)]
    for n in ("expand_bits", "contract_bits"):
        f = getattr(L, "ogmpm_" + n)
        f.argtypes = [C.c_uint32]
        f.restype = C.c_uint32
    L.ogmpm_morton_code.argtypes = [C.c_uint32] * 3
    L.ogmpm_morton_code.restype = C.c_uint32
    L.ogmpm_cell_index.argtypes = [C.c_uint32] * 3
    L.ogmpm_cell_index.restype = C.c_uint32
    L.ogmpm_inverse_cell_index.argtypes = [C.c_uint32, _U32P]
    L.ogmpm_stable_sort_low_bits.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
    for suf, ct in (("f32", C.c_float), ("f64", C.c_double)):
        g = lambda n: getattr(L, "ogmpm_%s_%s" % (n, suf))
        g("create").argtypes = [C.POINTER(OracleParams)]
        g("create").restype = C.c_void_p
        g("destroy").argtypes = [C.c_void_p]
        g("add_qr_cloth").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        g("add_particles").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int] + [C.c_double] * 4
        g("migrate_out").argtypes = [C.c_void_p, C.c_int, ct, ct, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        g("migrate_out").restype = C.c_size_t
        g("migrate_in").argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        g("n_volumetric").argtypes = [C.c_void_p]
        g("n_volumetric").restype = C.c_size_t
        g("svd3").argtypes = [C.c_void_p] * 4
        g("volumetric_update").argtypes = [C.c_int] + [C.c_double] * 4 + [C.c_void_p, ct, C.c_void_p, C.c_void_p]
        g("finalize").argtypes = [C.c_void_p]
        g("rebuild_mapping").argtypes = [C.c_void_p, C.c_int]
        g("calc_fem").argtypes = [C.c_void_p, ct]
        g("p2g").argtypes = [C.c_void_p, ct]
        g("update_grid").argtypes = [C.c_void_p, C.c_int]
        g("g2p").argtypes = [C.c_void_p, ct]
        g("set_contacts").argtypes = [C.c_void_p, C.c_size_t] + [C.c_void_p] * 7
        g("update_contact").argtypes = [C.c_void_p, ct, ct, ct, ct, C.c_int, C.c_int, C.POINTER(ContactStats)]
        g("set_num_bodies").argtypes = [C.c_void_p, C.c_size_t]
        g("body_impulses").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        g("dump_cpu_state").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        g("contact_grad_hess").argtypes = [C.POINTER(OracleParams), ct, ct, ct, ct, ct] + [C.c_void_p] * 4
        g("contact_cost").argtypes = [C.POINTER(OracleParams), ct, ct, ct, ct, ct, C.c_void_p, C.c_void_p]
        g("contact_cost").restype = ct
        g("make_from_one_unit_vector").argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        g("compute_dphi_dF").argtypes = [C.POINTER(OracleParams), C.c_void_p, C.c_void_p]
        g("project_strain").argtypes = [C.POINTER(OracleParams), C.c_void_p]
        g("givens_qr33").argtypes = [C.c_void_p] * 3
        g("svd2x2").argtypes = [C.c_void_p] * 4
        for n in ("n_verts", "n_faces", "n_particles", "n_cells", "n_contacts"):
            g(n).argtypes = [C.c_void_p]
            g(n).restype = C.c_size_t
        g("touched_count").argtypes = [C.c_void_p]
        g("touched_count").restype = C.c_uint32
        g("view").argtypes = [C.c_void_p, C.c_char_p]
        g("view").restype = C.c_void_p
        g("substep_omp").argtypes = [C.c_void_p, ct, C.c_int, C.c_int]


def default_params(**kw):
    p = OracleParams()
    lib().ogmpm_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


# name -> (kind, components) ; kind: 'T' scalar, 'i' int32, 'u' uint32 ; sized by N/F/cells/blocks/Nc
_VIEWS = {
    "positions": ("T", "N", 3), "velocities": ("T", "N", 3), "volumes": ("T", "N", 1),
    "affine_matrices": ("T", "N", 9), "pids": ("i", "N", 1), "sort_keys": ("u", "N", 1),
    "sort_ids": ("u", "N", 1), "forces": ("T", "N", 3), "taus": ("T", "N", 9),
    "index_mappings": ("i", "N", 1), "indices": ("i", "F", 3), "Dm_inverses": ("T", "F", 4),
    "deformation_gradients": ("T", "F", 9), "grid_masses": ("T", "cells", 1),
    "grid_momentum": ("T", "cells", 3), "grid_v_star": ("T", "cells", 3),
    "grid_touched_flags": ("u", "blocks", 1), "grid_touched_ids": ("u", "blocks", 1),
    "grid_Hess": ("T", "cells", 9), "grid_Grad": ("T", "cells", 3), "grid_Dir": ("T", "cells", 3),
    "particle_F": ("T", "Nv", 9), "contact_vel": ("T", "Nc", 3), "contact_vel0": ("T", "Nc", 3), "contact_sort_keys": ("u", "Nc", 1),
}


class OracleMpm:
    """Same call sequence as the reference's GpuMpmState/GpuMpmSolver, on the CPU restatement."""

    def __init__(self, dtype=np.float64, params=None, fast=False):
        """fast=True: the timing-only `-O3 -march=native` build (bench.py only; never a parity checker)."""
        self.dtype = np.dtype(dtype)
        self.suf = "f32" if self.dtype == np.float32 else "f64"
        self.ct = C.c_float if self.suf == "f32" else C.c_double
        self.params = params if params is not None else default_params()
        self._L = lib(fast)
        self.h = self._f("create")(C.byref(self.params))

    def _f(self, name):
        return getattr(self._L, "ogmpm_%s_%s" % (name, self.suf))

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- GpuMpmState
    def AddQRCloth(self, pos, vel, indices):
        pos = np.ascontiguousarray(pos, self.dtype).reshape(-1, 3)
        vel = np.ascontiguousarray(vel, self.dtype).reshape(-1, 3)
        tri = np.ascontiguousarray(indices, np.int32).reshape(-1, 3)
        self._f("add_qr_cloth")(self.h, _ptr(pos), _ptr(vel), pos.shape[0], _ptr(tri), tri.shape[0])

    MODELS = {"fixed_corotated": 1, "stvk_hencky": 2, "drucker_prager": 3}

    def AddParticles(self, pos, vel, volume, model="fixed_corotated", youngs=1e5, poisson=0.3, friction_angle_deg=30.0,
                     cohesion=0.0):
        pos = np.ascontiguousarray(pos, self.dtype).reshape(-1, 3)
        vel = np.ascontiguousarray(vel, self.dtype).reshape(-1, 3)
        vol = np.ascontiguousarray(np.broadcast_to(np.asarray(volume, self.dtype), (pos.shape[0],)))
        self._f("add_particles")(self.h, _ptr(pos), _ptr(vel), _ptr(vol), pos.shape[0], self.MODELS[model],
                                 float(youngs), float(poisson), float(friction_angle_deg), float(cohesion))

    def migrate_out(self, axis, lo, hi, park, cap):
        rec = np.zeros((cap, 26), self.dtype)
        pids = np.zeros(cap, np.int32)
        park = np.ascontiguousarray(park, self.dtype)
        n = self._f("migrate_out")(self.h, int(axis), self.ct(lo), self.ct(hi), _ptr(park), _ptr(rec), _ptr(pids), cap)
        return rec[:n].copy(), pids[:n].copy()

    def migrate_count(self, axis, lo, hi, edges):
        """Dry run of migrate_out (restates migrate_count_kernel of drake_b200/csrc/migrate.cuh): leavers per
        destination slab (destination = number of edges <= x) and the number of inactive slots."""
        n_cloth = self.n_faces + self.n_verts
        volumetric = np.asarray(self.read("pids")) >= n_cloth
        x = np.asarray(self.read("positions")).reshape(-1, 3)[:, axis]
        vol = np.asarray(self.read("volumes"))
        lo, hi = self.dtype.type(lo), self.dtype.type(hi)
        leaves = volumetric & (vol > 0) & ~((x >= lo) & (x < hi))
        edges = np.asarray(edges, np.float64).reshape(-1)
        dest = np.searchsorted(edges, x[leaves].astype(np.float64), side="right")
        return np.bincount(dest, minlength=len(edges) + 1).astype(np.int64), int((volumetric & ~(vol > 0)).sum())

    def migrate_in(self, records):
        rec = np.ascontiguousarray(records, self.dtype).reshape(-1, 26)
        pids = np.zeros(rec.shape[0], np.int32)
        if rec.shape[0]:
            assert self._f("migrate_in")(self.h, _ptr(rec), rec.shape[0], _ptr(pids)) == 0, "no inactive particle left"
        return pids

    def Finalize(self):
        self._f("finalize")(self.h)

    @property
    def n_particles(self):
        return self._f("n_particles")(self.h)

    @property
    def n_faces(self):
        return self._f("n_faces")(self.h)

    @property
    def n_verts(self):
        return self._f("n_verts")(self.h)

    @property
    def n_contacts(self):
        return self._f("n_contacts")(self.h)

    def touched_count(self):
        return int(self._f("touched_count")(self.h))

    def read(self, name):
        kind, dim, comp = _VIEWS[name]
        n = {"N": self.n_particles, "F": self.n_faces, "cells": self._f("n_cells")(self.h),
             "blocks": self._f("n_cells")(self.h) // 64, "Nc": self.n_contacts,
             "Nv": self._f("n_volumetric")(self.h)}[dim]
        dt = {"T": self.dtype, "i": np.dtype(np.int32), "u": np.dtype(np.uint32)}[kind]
        p = self._f("view")(self.h, name.encode())
        if not p or n == 0:
            return np.zeros((0,), dt)
        buf = (C.c_char * (n * comp * dt.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dt).copy()

    def view(self, name):
        """Writable numpy array ALIASING the oracle's buffer (valid until the next call that reallocates)."""
        kind, dim, comp = _VIEWS[name]
        n = {"N": self.n_particles, "F": self.n_faces, "cells": self._f("n_cells")(self.h),
             "blocks": self._f("n_cells")(self.h) // 64, "Nc": self.n_contacts,
             "Nv": self._f("n_volumetric")(self.h)}[dim]
        dt = {"T": self.dtype, "i": np.dtype(np.int32), "u": np.dtype(np.uint32)}[kind]
        p = self._f("view")(self.h, name.encode())
        buf = (C.c_char * (n * comp * dt.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dt)

    # --- GpuMpmSolver
    def RebuildMapping(self, sort):
        self._f("rebuild_mapping")(self.h, 1 if sort else 0)

    def CalcFemStateAndForce(self, dt):
        self._f("calc_fem")(self.h, self.ct(dt))

    def ParticleToGrid(self, dt):
        self._f("p2g")(self.h, self.ct(dt))

    def UpdateGrid(self, mpm_bc=-1):
        self._f("update_grid")(self.h, int(mpm_bc))

    def GridToParticle(self, dt):
        self._f("g2p")(self.h, self.ct(dt))

    def CopyContactPairs(self, mpm_id, rigid_id, dist, normal, pos, rigid_v, rigid_p):
        a = lambda x, t: np.ascontiguousarray(x, t)
        mpm_id, rigid_id = a(mpm_id, np.uint32), a(rigid_id, np.uint32)
        dist, normal, pos = a(dist, self.dtype), a(normal, self.dtype), a(pos, self.dtype)
        rigid_v, rigid_p = a(rigid_v, self.dtype), a(rigid_p, self.dtype)
        self._f("set_contacts")(self.h, mpm_id.size, _ptr(mpm_id), _ptr(rigid_id), _ptr(dist), _ptr(normal),
                                _ptr(pos), _ptr(rigid_v), _ptr(rigid_p))

    def UpdateContact(self, dt, mu, stiffness, damping, exact_line_search=False, max_iterations=0):
        st = ContactStats()
        self._f("update_contact")(self.h, self.ct(dt), self.ct(mu), self.ct(stiffness), self.ct(damping),
                                  1 if exact_line_search else 0, int(max_iterations), C.byref(st))
        return st

    def ReallocateExternelBodies(self, n):
        self._nb = n
        self._f("set_num_bodies")(self.h, n)

    def ExternelBodyForceToHost(self):
        tau = np.zeros((self._nb, 3), self.dtype)
        f = np.zeros((self._nb, 3), self.dtype)
        self._f("body_impulses")(self.h, _ptr(tau), _ptr(f))
        return tau, f

    def DumpCpuState(self):
        pos = np.zeros((self.n_verts, 3), self.dtype)
        tri = np.zeros((self.n_faces, 3), np.int32)
        self._f("dump_cpu_state")(self.h, _ptr(pos), _ptr(tri))
        return pos, tri

    def substep(self, dt, mpm_bc=-1, sort=False):
        self.RebuildMapping(sort)
        self.CalcFemStateAndForce(dt)
        self.ParticleToGrid(dt)
        self.UpdateGrid(mpm_bc)
        self.GridToParticle(dt)

    def substep_omp(self, dt, mpm_bc=-1, threads=1):
        self._f("substep_omp")(self.h, self.ct(dt), int(mpm_bc), int(threads))
