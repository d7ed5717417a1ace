// gmpm_capi.cu -- the extern "C" boundary of libgmpm_b200.so (include/gmpm_c.h).
#include <mutex>
#include <new>
#include <string>
#define GMPM_STR2(x) #x
#define GMPM_STR(x) GMPM_STR2(x)

#include "contact.cuh"
#include "checkpoint.cuh"
#include "contact_gen.cuh"
#include "getters.cuh"
#include "migrate.cuh"
#include "state.cuh"

namespace {
thread_local std::string g_last_error;

template <typename Fn>
int guarded(Fn&& fn) {
  try {
    fn();
    return GMPM_OK;
  } catch (const gmpm::CudaError& e) {
    g_last_error = e.what();
    if (e.code == cudaErrorNoDevice || e.code == cudaErrorInsufficientDriver) return GMPM_ERR_NO_DEVICE;
    if (e.code == cudaErrorMemoryAllocation) return GMPM_ERR_OOM;
    return GMPM_ERR_CUDA;
  } catch (const std::out_of_range& e) {
    g_last_error = e.what();
    return GMPM_ERR_UNKNOWN_NAME;
  } catch (const std::bad_alloc& e) {
    g_last_error = e.what();
    return GMPM_ERR_OOM;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return GMPM_ERR_INVALID;
  }
}
}  // namespace

extern "C" {

void gmpm_default_params(gmpm_params* p) {
  p->domain_bits = 7;
  p->boundary_cells = 3;
  p->gravity_axis = 2;
  p->device = -1;
  p->youngs = 400000.;
  p->poisson = .3;
  p->density = 2000.;
  p->gamma = 0.;
  p->k_normal = 100000.;
  p->v_blend = .8;
  p->c_f = 0.;
  p->sdf_friction = 0.3;
  p->gravity = -9.8;
  p->epsv = 1e-3;
  p->dx = 0.;
}
const char* gmpm_last_error(void) { return g_last_error.c_str(); }
const char* gmpm_version(void) { return "gmpm_b200 0.1 (sm_100a, CUDA " GMPM_STR(CUDART_VERSION) ")"; }
int gmpm_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
int gmpm_device_sync(void) {
  return guarded([&] { GMPM_CUDA(cudaDeviceSynchronize()); });
}
uint32_t gmpm_morton_code(uint32_t x, uint32_t y, uint32_t z) { return gmpm::morton3(x, y, z); }
uint32_t gmpm_cell_index(uint32_t x, uint32_t y, uint32_t z) { return gmpm::cell_key(x, y, z); }
void gmpm_inverse_cell_index(uint32_t index, uint32_t* xyz) { gmpm::inverse_cell_key(index, xyz[0], xyz[1], xyz[2]); }

int gmpm_radix_sort_pairs(const uint32_t* keys_in, const uint32_t* ids_in, size_t n, int num_bit, uint32_t* keys_out,
                          uint32_t* ids_out) {
  return guarded([&] {
    if (num_bit < 0 || num_bit > 32) throw std::invalid_argument("num_bit out of range");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
      throw gmpm::CudaError(cudaErrorNoDevice, "no CUDA device available (libgmpm_b200 has no CPU fallback)");
    if (n == 0) return;
    uint32_t *k[2], *v[2], *kin, *vin, *scratch;
    for (int i = 0; i < 2; ++i) {
      GMPM_CUDA(cudaMalloc(&k[i], n * 4));
      GMPM_CUDA(cudaMalloc(&v[i], n * 4));
    }
    GMPM_CUDA(cudaMalloc(&kin, n * 4));
    GMPM_CUDA(cudaMalloc(&vin, n * 4));
    GMPM_CUDA(cudaMalloc(&scratch, gmpm::radix_sort_scratch_words(n) * 4));
    GMPM_CUDA(cudaMemcpy(kin, keys_in, n * 4, cudaMemcpyHostToDevice));
    GMPM_CUDA(cudaMemcpy(vin, ids_in, n * 4, cudaMemcpyHostToDevice));
    int r = 0;
    if (num_bit == 0) {
      GMPM_CUDA(cudaMemcpy(k[0], kin, n * 4, cudaMemcpyDeviceToDevice));
      GMPM_CUDA(cudaMemcpy(v[0], vin, n * 4, cudaMemcpyDeviceToDevice));
    } else {
      int dev = 0, sms = 0;
      GMPM_CUDA(cudaGetDevice(&dev));
      GMPM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      gmpm::SortArgs a = gmpm::make_sort_args(kin, vin, k, v, (uint32_t)n, num_bit, scratch);
      const int ctas = gmpm::sort_grid_ctas(gmpm::radix_sort_kernel, sms, a.n_tiles);
      void* kargs[] = {&a};
      GMPM_CUDA(cudaLaunchCooperativeKernel((void*)gmpm::radix_sort_kernel, dim3(ctas), dim3(gmpm::kSortThreads), kargs,
                                            gmpm::kSortSmemBytes, 0));
      r = (a.n_passes - 1) & 1;
    }
    GMPM_CUDA(cudaDeviceSynchronize());
    GMPM_CUDA(cudaMemcpy(keys_out, k[r], n * 4, cudaMemcpyDeviceToHost));
    GMPM_CUDA(cudaMemcpy(ids_out, v[r], n * 4, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 2; ++i) {
      cudaFree(k[i]);
      cudaFree(v[i]);
    }
    cudaFree(kin);
    cudaFree(vin);
    cudaFree(scratch);
  });
}

// ---------------------------------------------------------------------------------------------------
// gmpm_radix_sort_pairs_static_scratch: the same sort for callers that must not allocate (see gmpm_c.h).
// Everything the sort needs beyond the caller's input and output arrays lives in ONE array that is part
// of the library image: the driver places it on every GPU that loads the library and keeps it there
// until the process exits.  Layout (uint32 words), sized for kStaticSortMaxPairs pairs:
//   [0, M)        the "other" key buffer of the ping-pong (the caller's keys_out is the final one)
//   [M, 2M)       the "other" id buffer
//   [2M, 2M + H)  per-tile digit histograms and digit totals (H = radix_sort_scratch_words(M))
constexpr size_t kStaticSortMaxPairs = GMPM_RADIX_SORT_STATIC_MAX_PAIRS;  // gmpm_c.h
constexpr size_t kStaticSortHistWords =
    (size_t)gmpm::kMaxRadix * ((kStaticSortMaxPairs + gmpm::kSortWarps * 32 * gmpm::kSortItemsSmall - 1) /
                               (gmpm::kSortWarps * 32 * gmpm::kSortItemsSmall)) + gmpm::kMaxRadix;
constexpr size_t kStaticSortWords = 2 * kStaticSortMaxPairs + kStaticSortHistWords;
static __device__ __align__(256) uint32_t g_static_sort_scratch[kStaticSortWords];

namespace {
constexpr int kStaticSortMaxDevices = 64;
std::mutex g_static_sort_mu;
uint32_t* g_static_sort_ptr[kStaticSortMaxDevices] = {};  // device address of the array, per device ordinal
int g_static_sort_ctas[kStaticSortMaxDevices] = {};       // co-resident CTAs of radix_sort_kernel, per device

// Set-up of the CURRENT device, done by the first call that runs on it.  In host code the name
// g_static_sort_scratch is only a handle for the runtime, not somewhere a kernel could read or write:
// the address the kernel gets is whatever cudaGetSymbolAddress reports for the device that is current.
void static_sort_setup(int dev, uint32_t** scratch, int* max_ctas) {
  if (dev < 0 || dev >= kStaticSortMaxDevices) throw std::invalid_argument("device ordinal beyond the per-device table");
  std::lock_guard<std::mutex> lock(g_static_sort_mu);
  if (!g_static_sort_ptr[dev]) {
    uint32_t* ptr = nullptr;
    GMPM_CUDA(cudaGetSymbolAddress((void**)&ptr, g_static_sort_scratch));
    if (!ptr) throw gmpm::CudaError(cudaErrorInvalidSymbol, "cudaGetSymbolAddress returned a null scratch address");
    int sms = 0, occ = 0;
    GMPM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    GMPM_CUDA(cudaFuncSetAttribute(gmpm::radix_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)gmpm::kSortSmemBytes));
    GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gmpm::radix_sort_kernel, gmpm::kSortThreads,
                                                            gmpm::kSortSmemBytes));
    if (occ < 1) throw gmpm::CudaError(cudaErrorLaunchOutOfResources, "radix sort kernel does not fit on an SM");
    g_static_sort_ctas[dev] = sms * occ;
    g_static_sort_ptr[dev] = ptr;
  }
  *scratch = g_static_sort_ptr[dev];
  *max_ctas = g_static_sort_ctas[dev];
}

// true if the pointer is one a kernel running on the current device may dereference
bool device_readable(const void* p) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}
bool ranges_overlap(const void* a, const void* b, size_t bytes) {
  const uintptr_t x = (uintptr_t)a, y = (uintptr_t)b;
  return x < y + bytes && y < x + bytes;
}
}  // namespace

int gmpm_radix_sort_pairs_static_scratch(const uint32_t* keys_in_dev, const uint32_t* ids_in_dev, size_t n, int num_bit,
                                         uint32_t* keys_out_dev, uint32_t* ids_out_dev, void* cuda_stream) {
  return guarded([&] {
    if (num_bit < 0 || num_bit > 32) throw std::invalid_argument("num_bit out of range");
    if (n > kStaticSortMaxPairs)
      throw std::invalid_argument("gmpm_radix_sort_pairs_static_scratch: n = " + std::to_string(n) +
                                  " pairs is more than the built-in scratch array holds (" +
                                  std::to_string(kStaticSortMaxPairs) +
                                  "); use gmpm_radix_sort_pairs, which allocates, or split the input");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
      throw gmpm::CudaError(cudaErrorNoDevice, "no CUDA device available (libgmpm_b200 has no CPU fallback)");
    if (n == 0) return;
    if (!keys_in_dev || !ids_in_dev || !keys_out_dev || !ids_out_dev) throw std::invalid_argument("null array pointer");
    const void* arrays[4] = {keys_in_dev, ids_in_dev, keys_out_dev, ids_out_dev};
    for (int i = 0; i < 4; ++i) {
      if (!device_readable(arrays[i]))
        throw std::invalid_argument("gmpm_radix_sort_pairs_static_scratch takes DEVICE arrays (argument " +
                                    std::to_string(i) + " is not device memory); it cannot stage host arrays "
                                    "because it allocates nothing");
      for (int j = i + 1; j < 4; ++j)
        if (ranges_overlap(arrays[i], arrays[j], n * 4))
          throw std::invalid_argument("input and output arrays must not overlap (the sort is not in place)");
    }
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    int dev = 0, max_ctas = 0;
    uint32_t* scratch = nullptr;
    GMPM_CUDA(cudaGetDevice(&dev));
    static_sort_setup(dev, &scratch, &max_ctas);
    if (num_bit == 0) {  // nothing to sort on: the pairs pass through in input order
      GMPM_CUDA(cudaMemcpyAsync(keys_out_dev, keys_in_dev, n * 4, cudaMemcpyDeviceToDevice, stream));
      GMPM_CUDA(cudaMemcpyAsync(ids_out_dev, ids_in_dev, n * 4, cudaMemcpyDeviceToDevice, stream));
      return;
    }
    // pass p writes buffer pair p & 1, so the last of n_passes passes writes pair (n_passes - 1) & 1: that pair is
    // the caller's output, the other one is the built-in array.  The input arrays are only ever read.
    gmpm::SortPass plan[4];
    const int last = (gmpm::plan_sort_passes(num_bit, plan) - 1) & 1;
    uint32_t *k[2], *v[2];
    k[last] = keys_out_dev;
    v[last] = ids_out_dev;
    k[last ^ 1] = scratch;
    v[last ^ 1] = scratch + kStaticSortMaxPairs;
    gmpm::SortArgs a =
        gmpm::make_sort_args(keys_in_dev, ids_in_dev, k, v, (uint32_t)n, num_bit, scratch + 2 * kStaticSortMaxPairs);
    // Nothing in the array has to start at zero: every histogram row and digit total a pass reads was written by
    // that same pass, and the ping-pong half is written before it is read, so what the previous call left is dead.
    const int want = a.n_tiles > 64u ? (int)a.n_tiles : 64;  // same grid as sort_grid_ctas
    const int ctas = want < max_ctas ? want : max_ctas;
    void* kargs[] = {&a};
    GMPM_CUDA(cudaLaunchCooperativeKernel((void*)gmpm::radix_sort_kernel, dim3(ctas), dim3(gmpm::kSortThreads), kargs,
                                          gmpm::kSortSmemBytes, stream));
  });
}

#define GMPM_DEFINE(SUF, T)                                                                                    \
  struct gmpm_state_##SUF {                                                                                    \
    gmpm::State<T> st;                                                                                         \
  };                                                                                                           \
  int gmpm_create_##SUF(const gmpm_params* p, gmpm_state_##SUF** out) {                                        \
    return guarded([&] {                                                                                       \
      if (!out) throw std::invalid_argument("null out pointer");                                               \
      *out = nullptr;                                                                                          \
      gmpm_params prm;                                                                                         \
      if (p) prm = *p; else gmpm_default_params(&prm);                                                         \
      auto* s = new gmpm_state_##SUF();                                                                        \
      try {                                                                                                    \
        s->st.init(prm);                                                                                       \
      } catch (...) {                                                                                          \
        delete s;                                                                                              \
        throw;                                                                                                 \
      }                                                                                                        \
      *out = s;                                                                                                \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_destroy_##SUF(gmpm_state_##SUF* s) {                                                                \
    return guarded([&] {                                                                                       \
      if (!s) return;                                                                                          \
      s->st.destroy();                                                                                         \
      delete s;                                                                                                \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_add_qr_cloth_##SUF(gmpm_state_##SUF* s, const T* pos, const T* vel, size_t nv, const int* tri,       \
                              size_t nt) {                                                                     \
    return guarded([&] { s->st.add_qr_cloth(pos, vel, nv, tri, nt); });                                        \
  }                                                                                                            \
  int gmpm_add_particles_##SUF(gmpm_state_##SUF* s, const T* pos, const T* vel, const T* volume, size_t n,     \
                               const gmpm_material* mat) {                                                     \
    return guarded([&] { s->st.add_particles(pos, vel, volume, n, mat); });                                    \
  }                                                                                                            \
  int gmpm_migrate_out_##SUF(gmpm_state_##SUF* s, int axis, T lo, T hi, const T* park, T* records, int* pids,  \
                             size_t cap, size_t* n) {                                                          \
    return guarded([&] {                                                                                       \
      const size_t c = s->st.migrate_out(axis, lo, hi, park, records, pids, cap);                              \
      if (n) *n = c;                                                                                           \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_migrate_count_##SUF(gmpm_state_##SUF* s, int axis, T lo, T hi, const double* edges, size_t n_edges, \
                               size_t* leavers, size_t* n_free) {                                              \
    return guarded([&] { s->st.migrate_count(axis, lo, hi, edges, n_edges, leavers, n_free); });               \
  }                                                                                                            \
  int gmpm_migrate_in_##SUF(gmpm_state_##SUF* s, const T* records, size_t n, int* pids) {                      \
    return guarded([&] { s->st.migrate_in(records, n, pids); });                                               \
  }                                                                                                            \
  int gmpm_finalize_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.finalize(); }); }                  \
  int gmpm_rebuild_mapping_##SUF(gmpm_state_##SUF* s, int sort) {                                              \
    return guarded([&] { s->st.rebuild_mapping(sort != 0); });                                                 \
  }                                                                                                            \
  int gmpm_calc_fem_##SUF(gmpm_state_##SUF* s, T dt) { return guarded([&] { s->st.calc_fem(dt); }); }          \
  int gmpm_p2g_##SUF(gmpm_state_##SUF* s, T dt) { return guarded([&] { s->st.p2g(dt); }); }                    \
  int gmpm_update_grid_##SUF(gmpm_state_##SUF* s, int bc) { return guarded([&] { s->st.update_grid(bc); }); }  \
  int gmpm_g2p_##SUF(gmpm_state_##SUF* s, T dt) { return guarded([&] { s->st.g2p(dt); }); }                    \
  int gmpm_sync_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.sync(); }); }                          \
  int gmpm_sync_particle_state_to_cpu_##SUF(gmpm_state_##SUF* s, T* out) {                                     \
    return guarded([&] { s->st.positions_to_host(out); });                                                     \
  }                                                                                                            \
  int gmpm_sync_particle_state_to_cpu_async_##SUF(gmpm_state_##SUF* s) {                                       \
    return guarded([&] { s->st.positions_async(); });                                                          \
  }                                                                                                            \
  int gmpm_set_query_bounds_##SUF(gmpm_state_##SUF* s, const T* lo_hi, size_t n) {                             \
    return guarded([&] { s->st.set_query_bounds(lo_hi, n); });                                                 \
  }                                                                                                            \
  int gmpm_clear_query_bounds_##SUF(gmpm_state_##SUF* s) {                                                     \
    return guarded([&] { s->st.clear_query_bounds(); });                                                       \
  }                                                                                                            \
  int gmpm_snapshot_d2h_bytes_##SUF(gmpm_state_##SUF* s, uint64_t* bytes) {                                    \
    return guarded([&] { if (bytes) *bytes = s->st.snapshot_d2h_bytes; });                                     \
  }                                                                                                            \
  int gmpm_positions_host_##SUF(gmpm_state_##SUF* s, const T** out) {                                          \
    return guarded([&] {                                                                                       \
      if (!out) throw std::invalid_argument("null out pointer");                                               \
      *out = s->st.wait_positions();                                                                           \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_copy_contact_pairs_##SUF(gmpm_state_##SUF* s, size_t n, const uint32_t* mpm_id,                     \
                                    const uint32_t* rigid_id, const T* dist, const T* normal, const T* pos,    \
                                    const T* rigid_v, const T* rigid_p) {                                      \
    return guarded([&] { s->st.copy_contact_pairs(n, mpm_id, rigid_id, dist, normal, pos, rigid_v, rigid_p); });\
  }                                                                                                            \
  int gmpm_generate_contact_pairs_##SUF(gmpm_state_##SUF* s, const gmpm_analytic_geometry* geoms, size_t n_geoms,\
                                        size_t* n_pairs) {                                                     \
    return guarded([&] {                                                                                       \
      const size_t n = s->st.generate_contact_pairs(geoms, n_geoms);                                           \
      if (n_pairs) *n_pairs = n;                                                                               \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_run_driver_substeps_##SUF(gmpm_state_##SUF* s, int n, T dt, int bc, int sort_every,                 \
                                     const gmpm_analytic_geometry* geoms, size_t n_geoms, int query_every, T mu,\
                                     T stiffness, T damping, int exact, int max_it, gmpm_driver_stats* stats) { \
    return guarded([&] {                                                                                       \
      s->st.run_driver_substeps(n, dt, bc, sort_every, geoms, n_geoms, query_every, mu, stiffness, damping,    \
                                exact != 0, max_it, stats);                                                    \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_set_driver_geometry_frozen_##SUF(gmpm_state_##SUF* s, int frozen) {                                 \
    return guarded([&] { s->st.driver_geometry_frozen = frozen != 0; });                                       \
  }                                                                                                            \
  int gmpm_update_contact_##SUF(gmpm_state_##SUF* s, T dt, T mu, T stiffness, T damping, int exact, int max_it,\
                                gmpm_contact_stats* stats) {                                                   \
    return guarded([&] { s->st.update_contact(dt, mu, stiffness, damping, exact != 0, max_it, stats); });      \
  }                                                                                                            \
  int gmpm_reallocate_external_bodies_##SUF(gmpm_state_##SUF* s, size_t n) {                                   \
    return guarded([&] { s->st.reallocate_external_bodies(n); });                                              \
  }                                                                                                            \
  int gmpm_external_body_force_to_host_##SUF(gmpm_state_##SUF* s, T* tau, T* f) {                              \
    return guarded([&] { s->st.external_body_force_to_host(tau, f); });                                        \
  }                                                                                                            \
  int gmpm_dump_cpu_state_##SUF(gmpm_state_##SUF* s, T* vert_pos, int* tri) {                                  \
    return guarded([&] { s->st.dump_cpu_state(vert_pos, tri); });                                              \
  }                                                                                                            \
  size_t gmpm_n_verts_##SUF(const gmpm_state_##SUF* s) { return s->st.n_verts; }                               \
  size_t gmpm_n_faces_##SUF(const gmpm_state_##SUF* s) { return s->st.n_faces; }                               \
  size_t gmpm_n_particles_##SUF(const gmpm_state_##SUF* s) { return s->st.N; }                                 \
  size_t gmpm_n_volumetric_##SUF(const gmpm_state_##SUF* s) { return s->st.n_vol; }                            \
  size_t gmpm_n_cells_##SUF(const gmpm_state_##SUF* s) { return s->st.k.n_cells; }                             \
  size_t gmpm_num_contacts_##SUF(const gmpm_state_##SUF* s) { return s->st.n_contacts; }                       \
  int gmpm_grid_touched_cnt_host_##SUF(gmpm_state_##SUF* s, uint32_t* out) {                                   \
    return guarded([&] { *out = s->st.touched_count_host(); });                                                \
  }                                                                                                            \
  int gmpm_read_##SUF(gmpm_state_##SUF* s, const char* name, void* out, size_t cap, size_t* written) {         \
    return guarded([&] {                                                                                       \
      const size_t b = s->st.read(name, out, cap);                                                             \
      if (written) *written = b;                                                                               \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_flags_##SUF(gmpm_state_##SUF* s, void** flags_dev, size_t* n_blocks) {                         \
    return guarded([&] {                                                                                       \
      if (!flags_dev || !n_blocks) throw std::invalid_argument("null out pointer");                            \
      *flags_dev = s->st.halo_flags();                                                                         \
      *n_blocks = s->st.k.n_blocks;                                                                            \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_plan_async_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.halo_plan_async(); }); }      \
  int gmpm_halo_reset_plan_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.halo_reset_plan(); }); }      \
  int gmpm_halo_pack_##SUF(gmpm_state_##SUF* s, uint32_t* n_shared, void** packed_dev) {                       \
    return guarded([&] {                                                                                       \
      if (!n_shared || !packed_dev) throw std::invalid_argument("null out pointer");                           \
      *n_shared = s->st.halo_pack();                                                                           \
      *packed_dev = s->st.halo_packed;                                                                         \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_unpack_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.halo_unpack(); }); }            \
  int gmpm_halo_peer_init_##SUF(gmpm_state_##SUF* s, uint32_t rank, uint32_t world, uint32_t cap_blocks) {     \
    return guarded([&] { s->st.halo_peer_init(rank, world, cap_blocks); });                                    \
  }                                                                                                            \
  int gmpm_halo_peer_window_##SUF(gmpm_state_##SUF* s, unsigned char* ipc_handle64, void** local_ptr) {        \
    return guarded([&] {                                                                                       \
      if (ipc_handle64) s->st.halo_peer_ipc_handle(ipc_handle64);                                              \
      s->st.require_peer("halo_peer_window");                                                                  \
      if (local_ptr) *local_ptr = s->st.halo_window;                                                           \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_peer_attach_##SUF(gmpm_state_##SUF* s, uint32_t peer, const unsigned char* ipc_handle64,       \
                                  void* same_process_ptr) {                                                    \
    return guarded([&] { s->st.halo_peer_attach(peer, ipc_handle64, same_process_ptr); });                     \
  }                                                                                                            \
  int gmpm_halo_peer_send_##SUF(gmpm_state_##SUF* s, uint32_t* n_shared) {                                     \
    return guarded([&] {                                                                                       \
      const uint32_t n = s->st.halo_peer_send();                                                               \
      if (n_shared) *n_shared = n;                                                                             \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_peer_recv_##SUF(gmpm_state_##SUF* s) { return guarded([&] { s->st.halo_peer_recv(); }); }      \
  int gmpm_update_grid_part_##SUF(gmpm_state_##SUF* s, int bc, int part) {                                     \
    return guarded([&] { s->st.update_grid_part(bc, part); });                                                 \
  }                                                                                                            \
  int gmpm_g2p_part_##SUF(gmpm_state_##SUF* s, T dt, int part) {                                               \
    return guarded([&] { s->st.g2p_part(dt, part); });                                                         \
  }                                                                                                            \
  int gmpm_stream_##SUF(gmpm_state_##SUF* s, void** cuda_stream) {                                             \
    return guarded([&] {                                                                                       \
      if (!cuda_stream) throw std::invalid_argument("null out pointer");                                       \
      *cuda_stream = (void*)s->st.stream;                                                                      \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_halo_stream_##SUF(gmpm_state_##SUF* s, void** cuda_stream) {                                        \
    return guarded([&] {                                                                                       \
      if (!cuda_stream) throw std::invalid_argument("null out pointer");                                       \
      s->st.require_finalized();                                                                               \
      s->st.halo_init();                                                                                       \
      *cuda_stream = (void*)s->st.halo_stream;                                                                 \
    });                                                                                                        \
  }                                                                                                            \
  size_t gmpm_checkpoint_bytes_##SUF(const gmpm_state_##SUF* s) { return s->st.checkpoint_bytes(); }           \
  int gmpm_save_state_##SUF(gmpm_state_##SUF* s, void* out, size_t cap, size_t* written) {                     \
    return guarded([&] {                                                                                       \
      const size_t b = s->st.save_state(out, cap);                                                             \
      if (written) *written = b;                                                                               \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_load_state_##SUF(gmpm_state_##SUF* s, const void* in, size_t bytes) {                               \
    return guarded([&] { s->st.load_state(in, bytes); });                                                      \
  }                                                                                                            \
  int gmpm_run_substeps_##SUF(gmpm_state_##SUF* s, int n, T dt, int bc, int sort_every) {                      \
    return guarded([&] { s->st.run_substeps(n, dt, bc, sort_every); });                                        \
  }                                                                                                            \
  uint64_t gmpm_kernel_launches_##SUF(const gmpm_state_##SUF* s) { return s->st.launches; }                    \
  int gmpm_set_profiling_##SUF(gmpm_state_##SUF* s, int on) {                                                  \
    return guarded([&] {                                                                                       \
      s->st.sync();                                                                                            \
      s->st.prof.reset();                                                                                      \
      s->st.prof.on = on != 0;                                                                                 \
    });                                                                                                        \
  }                                                                                                            \
  int gmpm_kernel_time_##SUF(gmpm_state_##SUF* s, const char* prefix, double* ms, uint64_t* launches) {        \
    return guarded([&] {                                                                                       \
      s->st.sync();                                                                                            \
      double t = 0;                                                                                            \
      uint64_t c = 0;                                                                                          \
      const std::string pre(prefix ? prefix : "");                                                             \
      for (auto& kv : s->st.prof.acc)                                                                          \
        if (kv.first.compare(0, pre.size(), pre) == 0) {                                                       \
          t += kv.second.first;                                                                                \
          c += kv.second.second;                                                                               \
        }                                                                                                      \
      if (ms) *ms = t;                                                                                         \
      if (launches) *launches = c;                                                                             \
    });                                                                                                        \
  }                                                                                                            \
  size_t gmpm_profile_report_##SUF(gmpm_state_##SUF* s, char* buf, size_t cap) {                               \
    std::string r;                                                                                             \
    if (guarded([&] {                                                                                          \
          s->st.sync();                                                                                        \
          char line[256];                                                                                      \
          for (auto& kv : s->st.prof.acc) {                                                                    \
            snprintf(line, sizeof(line), "%s %.6f %llu\n", kv.first.c_str(), kv.second.first,                  \
                     (unsigned long long)kv.second.second);                                                    \
            r += line;                                                                                         \
          }                                                                                                    \
        }) != GMPM_OK)                                                                                         \
      return 0;                                                                                                \
    if (buf && cap) {                                                                                          \
      const size_t n = r.size() < cap - 1 ? r.size() : cap - 1;                                                \
      memcpy(buf, r.data(), n);                                                                                \
      buf[n] = 0;                                                                                              \
    }                                                                                                          \
    return r.size() + 1;                                                                                       \
  }                                                                                                            \
  int gmpm_timer_mark_##SUF(gmpm_state_##SUF* s, int slot) {                                                   \
    return guarded([&] { s->st.timer_mark(slot); });                                                           \
  }                                                                                                            \
  int gmpm_timer_elapsed_ms_##SUF(gmpm_state_##SUF* s, int a, int b, double* ms) {                             \
    return guarded([&] { *ms = s->st.timer_elapsed_ms(a, b); });                                               \
  }                                                                                                            \
  size_t gmpm_device_bytes_##SUF(const gmpm_state_##SUF* s) { return s->st.device_bytes; }                     \
  int gmpm_block_counts_##SUF(gmpm_state_##SUF* s, uint32_t* pb, uint32_t* tb) {                               \
    return guarded([&] {                                                                                       \
      if (pb) *pb = s->st.pblock_count_host();                                                                 \
      if (tb) *tb = s->st.touched_count_host();                                                                \
    });                                                                                                        \
  }

GMPM_DEFINE(f32, float)
GMPM_DEFINE(f64, double)

}  // extern "C"
