"""ctypes binding of oracle/_ref/libgmpm_ref*.so: the REFERENCE's own multibody/gpu_mpm kernels compiled
for sm_100 from /root/reference by oracle/Makefile (fp32 + fp64 instantiations).  TEST INFRASTRUCTURE
ONLY (pins the CPU oracle on the GPU box; bench.py times it as the existing kernel to beat)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def available(fast=False):
    return os.path.exists(os.path.join(REF_DIR, "libgmpm_ref_fast.so" if fast else "libgmpm_ref.so"))


_libs = {}


def _lib(fast):
    if fast not in _libs:
        L = C.CDLL(os.path.join(REF_DIR, "libgmpm_ref_fast.so" if fast else "libgmpm_ref.so"))
        vp, sz = C.c_void_p, C.c_size_t
        for suf, ct in (("f32", C.c_float), ("f64", C.c_double)):
            g = lambda n: getattr(L, "rgmpm_%s_%s" % (n, suf))
            g("create").restype = vp
            g("destroy").argtypes = [vp]
            g("add_qr_cloth").argtypes = [vp, vp, vp, sz, vp, sz]
            g("finalize").argtypes = [vp]
            g("rebuild_mapping").argtypes = [vp, C.c_int]
            g("calc_fem").argtypes = [vp, ct]
            g("p2g").argtypes = [vp, ct]
            g("update_grid").argtypes = [vp, C.c_int]
            g("g2p").argtypes = [vp, ct]
            g("sync").argtypes = [vp]
            g("set_contacts").argtypes = [vp, sz] + [vp] * 7
            g("update_contact").argtypes = [vp, ct, ct, ct, ct, C.c_int]
            g("set_num_bodies").argtypes = [vp, sz]
            g("body_impulses").argtypes = [vp, vp, vp]
            g("touched_count").argtypes = [vp]
            g("touched_count").restype = C.c_uint32
            g("n_particles").argtypes = [vp]
            g("n_particles").restype = sz
            g("read").argtypes = [vp, C.c_char_p, vp, sz]
            g("read").restype = sz
        L.rgmpm_build_flags.restype = C.c_char_p
        _libs[fast] = L
    return _libs[fast]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class RefMpm:
    """The reference's GpuMpmState<T> + GpuMpmSolver<T> (fixed 128^3 grid, settings.h)."""

    def __init__(self, dtype=np.float64, fast=False):
        self.dtype = np.dtype(dtype)
        self.suf = "f32" if self.dtype == np.float32 else "f64"
        self.ct = C.c_float if self.suf == "f32" else C.c_double
        self._L = _lib(fast)
        self.h = self._f("create")()
        self.n_faces = 0
        self.n_verts = 0

    def _f(self, n):
        return getattr(self._L, "rgmpm_%s_%s" % (n, self.suf))

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def AddQRCloth(self, pos, vel, tri):
        pos = np.ascontiguousarray(pos, self.dtype).reshape(-1, 3)
        vel = np.ascontiguousarray(vel, self.dtype).reshape(-1, 3)
        tri = np.ascontiguousarray(tri, np.int32).reshape(-1, 3)
        self._f("add_qr_cloth")(self.h, _ptr(pos), _ptr(vel), pos.shape[0], _ptr(tri), tri.shape[0])
        self.n_verts += pos.shape[0]
        self.n_faces += tri.shape[0]

    def Finalize(self):
        self._f("finalize")(self.h)

    def RebuildMapping(self, sort):
        self._f("rebuild_mapping")(self.h, 1 if sort else 0)

    def CalcFemStateAndForce(self, dt):
        self._f("calc_fem")(self.h, self.ct(dt))

    def ParticleToGrid(self, dt):
        self._f("p2g")(self.h, self.ct(dt))

    def UpdateGrid(self, bc=-1):
        self._f("update_grid")(self.h, int(bc))

    def GridToParticle(self, dt):
        self._f("g2p")(self.h, self.ct(dt))

    def GpuSync(self):
        self._f("sync")(self.h)

    def substep(self, dt, bc=-1, sort=False):
        self.RebuildMapping(sort)
        self.CalcFemStateAndForce(dt)
        self.ParticleToGrid(dt)
        self.UpdateGrid(bc)
        self.GridToParticle(dt)

    def CopyContactPairs(self, mpm_id, rigid_id, dist, normal, pos, rigid_v, rigid_p):
        a = lambda x, t: np.ascontiguousarray(x, t)
        mpm_id, rigid_id = a(mpm_id, np.uint32), a(rigid_id, np.uint32)
        arrs = [a(x, self.dtype) for x in (dist, normal, pos, rigid_v, rigid_p)]
        self._f("set_contacts")(self.h, mpm_id.size, _ptr(mpm_id), _ptr(rigid_id), *[_ptr(x) for x in arrs])
        self._nc = mpm_id.size

    def UpdateContact(self, dt, mu, k, d, exact=False):
        self._f("update_contact")(self.h, self.ct(dt), self.ct(mu), self.ct(k), self.ct(d), 1 if exact else 0)

    def ReallocateExternelBodies(self, n):
        self._nb = n
        self._f("set_num_bodies")(self.h, n)

    def ExternelBodyForceToHost(self):
        tau, f = np.zeros((self._nb, 3), self.dtype), np.zeros((self._nb, 3), self.dtype)
        self._f("body_impulses")(self.h, _ptr(tau), _ptr(f))
        return tau, f

    def touched_count(self):
        return int(self._f("touched_count")(self.h))

    def read(self, name, count, np_dtype=None):
        dt = np.dtype(np_dtype) if np_dtype is not None else self.dtype
        out = np.zeros(count, dt)
        n = self._f("read")(self.h, name.encode(), _ptr(out), out.nbytes)
        assert n == out.nbytes, (name, n, out.nbytes)
        return out
