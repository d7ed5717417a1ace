"""ctypes binding of the C ABI (include/gmpm_c.h).  Fails loudly when the CUDA library is missing:
there is no CPU fallback in this package."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GMPM_B200_LIB") or os.path.join(HERE, "lib", "libgmpm_b200.so")  # override: build variants (tools/)


class GmpmError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("gmpm error %d: %s" % (status, msg))
        self.status = status


class Params(C.Structure):
    """Mirror of gmpm_params; defaults (gmpm_default_params) = reference settings.h:37-126."""
    _fields_ = [("domain_bits", C.c_int), ("boundary_cells", C.c_int), ("gravity_axis", C.c_int),
                ("device", C.c_int), ("youngs", C.c_double), ("poisson", C.c_double), ("density", C.c_double),
                ("gamma", C.c_double), ("k_normal", C.c_double), ("v_blend", C.c_double), ("c_f", C.c_double),
                ("sdf_friction", C.c_double), ("gravity", C.c_double), ("epsv", C.c_double), ("dx", C.c_double)]


class AnalyticGeometry(C.Structure):
    """Mirror of gmpm_analytic_geometry."""
    _fields_ = [("kind", C.c_int), ("rigid_id", C.c_uint32), ("lo", C.c_double * 3), ("hi", C.c_double * 3),
                ("velocity", C.c_double * 3), ("p_WB", C.c_double * 3)]


class Material(C.Structure):
    """Mirror of gmpm_material (volumetric materials, extension)."""
    _fields_ = [("model", C.c_int), ("reserved", C.c_int), ("youngs", C.c_double), ("poisson", C.c_double),
                ("friction_angle_deg", C.c_double), ("cohesion", C.c_double)]


MATERIAL_MODELS = {"fixed_corotated": 1, "stvk_hencky": 2, "drucker_prager": 3}


class DriverStats(C.Structure):
    """Mirror of gmpm_driver_stats."""
    _fields_ = [("substeps", C.c_int), ("contact_queries", C.c_int), ("newton_iterations", C.c_int),
                ("newton_iterations_max", C.c_int), ("contact_pairs", C.c_uint64), ("last_residual", C.c_double)]


class ContactStats(C.Structure):
    _fields_ = [("iterations", C.c_int), ("grid_dofs", C.c_int), ("line_search_evals", C.c_int),
                ("reserved", C.c_int), ("residual", C.c_double), ("energy", C.c_double)]


# every symbol include/gmpm_c.h declares (checked by tests/test_abi.py against the header text)
PLAIN_SYMBOLS = ["gmpm_default_params", "gmpm_last_error", "gmpm_version", "gmpm_device_count", "gmpm_device_sync",
                 "gmpm_morton_code", "gmpm_cell_index", "gmpm_inverse_cell_index", "gmpm_radix_sort_pairs",
                 "gmpm_radix_sort_pairs_static_scratch"]
TYPED_SYMBOLS = ["create", "destroy", "add_qr_cloth", "finalize", "rebuild_mapping", "calc_fem", "p2g",
                 "update_grid", "g2p", "sync", "sync_particle_state_to_cpu", "sync_particle_state_to_cpu_async",
                 "positions_host", "set_query_bounds", "add_particles", "migrate_out", "migrate_in", "migrate_count", "n_volumetric", "clear_query_bounds", "snapshot_d2h_bytes", "halo_flags", "halo_plan_async", "halo_reset_plan", "halo_pack", "halo_unpack", "halo_peer_init", "halo_peer_window", "halo_peer_attach", "halo_peer_send", "halo_peer_recv", "update_grid_part", "g2p_part", "stream", "halo_stream", "copy_contact_pairs",
                 "generate_contact_pairs", "run_driver_substeps", "set_driver_geometry_frozen", "update_contact", "reallocate_external_bodies", "external_body_force_to_host",
                 "dump_cpu_state", "n_verts", "n_faces", "n_particles", "n_cells", "num_contacts",
                 "grid_touched_cnt_host", "read", "checkpoint_bytes", "save_state", "load_state", "run_substeps", "kernel_launches", "set_profiling",
                 "kernel_time", "profile_report", "timer_mark", "timer_elapsed_ms", "device_bytes",
                 "block_counts"]

_lib = None


def lib():
    """Loads libgmpm_b200.so (never builds it implicitly on a GPU box: run `python -m drake_b200.build`
    or __graft_entry__.build() first)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `python -m drake_b200.build` (needs nvcc). "
                          "drake_b200 has no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    _declare(L)
    _lib = L
    return L


def _declare(L):
    vp, sz, u32p = C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32)
    L.gmpm_default_params.argtypes = [C.POINTER(Params)]
    L.gmpm_default_params.restype = None
    L.gmpm_last_error.restype = C.c_char_p
    L.gmpm_version.restype = C.c_char_p
    L.gmpm_device_count.restype = C.c_int
    L.gmpm_device_sync.restype = C.c_int
    for n in ("gmpm_morton_code", "gmpm_cell_index"):
        getattr(L, n).argtypes = [C.c_uint32] * 3
        getattr(L, n).restype = C.c_uint32
    L.gmpm_inverse_cell_index.argtypes = [C.c_uint32, u32p]
    L.gmpm_inverse_cell_index.restype = None
    L.gmpm_radix_sort_pairs.argtypes = [vp, vp, sz, C.c_int, vp, vp]
    L.gmpm_radix_sort_pairs_static_scratch.argtypes = [vp, vp, sz, C.c_int, vp, vp, vp]
    for suf, ct in (("f32", C.c_float), ("f64", C.c_double)):
        g = lambda n: getattr(L, "gmpm_%s_%s" % (n, suf))
        g("create").argtypes = [C.POINTER(Params), C.POINTER(vp)]
        g("destroy").argtypes = [vp]
        g("add_qr_cloth").argtypes = [vp, vp, vp, sz, vp, sz]
        g("add_particles").argtypes = [vp, vp, vp, vp, sz, C.POINTER(Material)]
        g("migrate_out").argtypes = [vp, C.c_int, ct, ct, vp, vp, vp, sz, C.POINTER(C.c_size_t)]
        g("migrate_in").argtypes = [vp, vp, sz, vp]
        g("migrate_count").argtypes = [vp, C.c_int, ct, ct, vp, sz, vp, C.POINTER(C.c_size_t)]
        g("n_volumetric").argtypes = [vp]
        g("n_volumetric").restype = sz
        g("finalize").argtypes = [vp]
        g("rebuild_mapping").argtypes = [vp, C.c_int]
        g("calc_fem").argtypes = [vp, ct]
        g("p2g").argtypes = [vp, ct]
        g("update_grid").argtypes = [vp, C.c_int]
        g("g2p").argtypes = [vp, ct]
        g("sync").argtypes = [vp]
        g("sync_particle_state_to_cpu").argtypes = [vp, vp]
        g("sync_particle_state_to_cpu_async").argtypes = [vp]
        g("positions_host").argtypes = [vp, C.POINTER(C.c_void_p)]
        g("checkpoint_bytes").argtypes = [vp]
        g("checkpoint_bytes").restype = sz
        g("save_state").argtypes = [vp, vp, sz, C.POINTER(sz)]
        g("load_state").argtypes = [vp, vp, sz]
        g("halo_flags").argtypes = [vp, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        g("halo_plan_async").argtypes = [vp]
        g("halo_reset_plan").argtypes = [vp]
        g("halo_pack").argtypes = [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_void_p)]
        g("halo_unpack").argtypes = [vp]
        g("halo_peer_init").argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32]
        g("halo_peer_window").argtypes = [vp, C.c_char_p, C.POINTER(C.c_void_p)]
        g("halo_peer_attach").argtypes = [vp, C.c_uint32, C.c_char_p, C.c_void_p]
        g("halo_peer_send").argtypes = [vp, C.POINTER(C.c_uint32)]
        g("halo_peer_recv").argtypes = [vp]
        g("update_grid_part").argtypes = [vp, C.c_int, C.c_int]
        g("g2p_part").argtypes = [vp, ct, C.c_int]
        g("stream").argtypes = [vp, C.POINTER(C.c_void_p)]
        g("halo_stream").argtypes = [vp, C.POINTER(C.c_void_p)]
        g("copy_contact_pairs").argtypes = [vp, sz] + [vp] * 7
        g("generate_contact_pairs").argtypes = [vp, C.POINTER(AnalyticGeometry), sz, C.POINTER(C.c_size_t)]
        g("update_contact").argtypes = [vp, ct, ct, ct, ct, C.c_int, C.c_int, C.POINTER(ContactStats)]
        g("run_driver_substeps").argtypes = [vp, C.c_int, ct, C.c_int, C.c_int, C.POINTER(AnalyticGeometry), sz, C.c_int,
                                             ct, ct, ct, C.c_int, C.c_int, C.POINTER(DriverStats)]
        g("set_driver_geometry_frozen").argtypes = [vp, C.c_int]
        g("set_query_bounds").argtypes = [vp, vp, sz]
        g("clear_query_bounds").argtypes = [vp]
        g("snapshot_d2h_bytes").argtypes = [vp, C.POINTER(C.c_uint64)]
        g("reallocate_external_bodies").argtypes = [vp, sz]
        g("external_body_force_to_host").argtypes = [vp, vp, vp]
        g("dump_cpu_state").argtypes = [vp, vp, vp]
        for n in ("n_verts", "n_faces", "n_particles", "n_cells", "num_contacts", "device_bytes"):
            g(n).argtypes = [vp]
            g(n).restype = sz
        g("grid_touched_cnt_host").argtypes = [vp, u32p]
        g("read").argtypes = [vp, C.c_char_p, vp, sz, C.POINTER(sz)]
        g("run_substeps").argtypes = [vp, C.c_int, ct, C.c_int, C.c_int]
        g("kernel_launches").argtypes = [vp]
        g("kernel_launches").restype = C.c_uint64
        g("set_profiling").argtypes = [vp, C.c_int]
        g("kernel_time").argtypes = [vp, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        g("profile_report").argtypes = [vp, C.c_char_p, sz]
        g("profile_report").restype = sz
        g("block_counts").argtypes = [vp, u32p, u32p]
        g("timer_mark").argtypes = [vp, C.c_int]
        g("timer_elapsed_ms").argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double)]


def check(status):
    if status != 0:
        raise GmpmError(status, lib().gmpm_last_error().decode())


def default_params(**kw):
    p = Params()
    lib().gmpm_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p
