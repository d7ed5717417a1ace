// ref_capi.cu -- thin extern "C" access to the UNMODIFIED reference classes
// drake::multibody::gmpm::GpuMpmState<T> / GpuMpmSolver<T> (compiled from /root/reference by
// oracle/Makefile into oracle/_ref/).  TEST INFRASTRUCTURE ONLY: used by tests/ (to pin the CPU
// oracle against the reference's own kernels on the B200) and by bench.py (as the "existing
// kernel to beat").  Nothing here is product code and the product never links it.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "multibody/gpu_mpm/cpu_mpm_model.h"
#include "multibody/gpu_mpm/cuda_mpm_model.cuh"
#include "multibody/gpu_mpm/cuda_mpm_solver.cuh"

namespace g = drake::multibody::gmpm;

template <typename T>
struct RefHandle {
  g::GpuMpmState<T> state;
  g::GpuMpmSolver<T> solver;
};

template <typename T>
static void ref_add_cloth(RefHandle<T>* h, const T* pos, const T* vel, size_t nv, const int* tri, size_t nf) {
  std::vector<g::Vec3<T>> p(nv), v(nv);
  for (size_t i = 0; i < nv; ++i) {
    p[i] = g::Vec3<T>(pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2]);
    v[i] = g::Vec3<T>(vel[i * 3], vel[i * 3 + 1], vel[i * 3 + 2]);
  }
  std::vector<int> idx(tri, tri + nf * 3);
  h->state.AddQRCloth(p, v, idx);
}

template <typename T>
static void ref_finalize(RefHandle<T>* h) {
  h->state.Finalize();
  // The reference never zero-initialises its grid (cuda_mpm_model.cu:98-103); do it through its
  // public accessors so that the first step is well defined.
  const size_t cells = g::config::G_DOMAIN_VOLUME, blocks = g::config::G_GRID_VOLUME;
  cudaMemset(h->state.grid_masses(), 0, cells * sizeof(T));
  cudaMemset(h->state.grid_momentum(), 0, cells * 3 * sizeof(T));
  cudaMemset(h->state.grid_v_star(), 0, cells * 3 * sizeof(T));
  cudaMemset(h->state.grid_touched_flags(), 0, blocks * sizeof(uint32_t));
  cudaMemset(h->state.grid_touched_ids(), 0, blocks * sizeof(uint32_t));
  cudaMemset(h->state.grid_Hess(), 0, cells * 9 * sizeof(T));
  cudaMemset(h->state.grid_Grad(), 0, cells * 3 * sizeof(T));
  cudaMemset(h->state.grid_Dir(), 0, cells * 3 * sizeof(T));
  cudaMemset(h->state.grid_alpha(), 0, cells * sizeof(T));
  cudaMemset(h->state.grid_E0(), 0, cells * sizeof(T));
  cudaMemset(h->state.grid_E1(), 0, cells * sizeof(T));
  cudaDeviceSynchronize();
}

template <typename T>
static void ref_set_contacts(RefHandle<T>* h, size_t n, const uint32_t* mpm_id, const uint32_t* rigid_id,
                             const T* dist, const T* normal, const T* pos, const T* rigid_v, const T* rigid_p) {
  g::MpmParticleContactPairs<T> pairs;
  for (size_t i = 0; i < n; ++i)
    pairs.push_back(mpm_id[i], rigid_id[i], dist[i], g::Vec3<T>(normal[i * 3], normal[i * 3 + 1], normal[i * 3 + 2]),
                    g::Vec3<T>(pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2]),
                    g::Vec3<T>(rigid_v[i * 3], rigid_v[i * 3 + 1], rigid_v[i * 3 + 2]),
                    g::Vec3<T>(rigid_p[i * 3], rigid_p[i * 3 + 1], rigid_p[i * 3 + 2]));
  h->solver.CopyContactPairs(&h->state, pairs);
}

// Copies a named device buffer to host.  Returns the number of BYTES written (0 = unknown name).
template <typename T>
static size_t ref_read(RefHandle<T>* h, const char* name, void* out, size_t cap) {
  auto& s = h->state;
  const size_t N = s.n_particles(), F = s.n_faces(), Nc = s.num_contacts();
  const size_t cells = g::config::G_DOMAIN_VOLUME, blocks = g::config::G_GRID_VOLUME;
  const void* src = nullptr;
  size_t bytes = 0;
#define RD(n, p, b) if (!strcmp(name, n)) { src = (p); bytes = (b); }
  RD("positions", s.current_positions(), N * 3 * sizeof(T))
  RD("velocities", s.current_velocities(), N * 3 * sizeof(T))
  RD("volumes", s.current_volumes(), N * sizeof(T))
  RD("affine_matrices", s.current_affine_matrices(), N * 9 * sizeof(T))
  RD("pids", s.current_pids(), N * sizeof(int))
  RD("sort_keys", s.current_sort_keys(), N * sizeof(uint32_t))
  RD("sort_ids", s.current_sort_ids(), N * sizeof(uint32_t))
  RD("forces", s.forces(), N * 3 * sizeof(T))
  RD("taus", s.taus(), N * 9 * sizeof(T))
  RD("index_mappings", s.index_mappings(), N * sizeof(int))
  RD("indices", s.indices(), F * 3 * sizeof(int))
  RD("Dm_inverses", s.Dm_inverses(), F * 4 * sizeof(T))
  RD("deformation_gradients", s.deformation_gradients(), F * 9 * sizeof(T))
  RD("grid_masses", s.grid_masses(), cells * sizeof(T))
  RD("grid_momentum", s.grid_momentum(), cells * 3 * sizeof(T))
  RD("grid_v_star", s.grid_v_star(), cells * 3 * sizeof(T))
  RD("grid_touched_flags", s.grid_touched_flags(), blocks * sizeof(uint32_t))
  RD("grid_touched_ids", s.grid_touched_ids(), blocks * sizeof(uint32_t))
  RD("grid_Hess", s.grid_Hess(), cells * 9 * sizeof(T))
  RD("grid_Grad", s.grid_Grad(), cells * 3 * sizeof(T))
  RD("grid_Dir", s.grid_Dir(), cells * 3 * sizeof(T))
  RD("contact_vel", s.contact_vel(), Nc * 3 * sizeof(T))
  RD("contact_vel0", s.contact_vel0(), Nc * 3 * sizeof(T))
  RD("contact_sort_keys", s.contact_sort_keys(), Nc * sizeof(uint32_t))
#undef RD
  if (!src || bytes > cap) return 0;
  cudaDeviceSynchronize();
  if (cudaMemcpy(out, src, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
  return bytes;
}

#define RGMPM_DEFINE(SUF, T)                                                                             \
  extern "C" void* rgmpm_create_##SUF() { return new RefHandle<T>(); }                                   \
  extern "C" void rgmpm_destroy_##SUF(void* h) {                                                         \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->state.Destroy();                                                                                  \
    delete r;                                                                                            \
  }                                                                                                      \
  extern "C" void rgmpm_add_qr_cloth_##SUF(void* h, const T* pos, const T* vel, size_t nv, const int* tri, \
                                           size_t nf) {                                                  \
    ref_add_cloth<T>(static_cast<RefHandle<T>*>(h), pos, vel, nv, tri, nf);                              \
  }                                                                                                      \
  extern "C" void rgmpm_finalize_##SUF(void* h) { ref_finalize<T>(static_cast<RefHandle<T>*>(h)); }      \
  extern "C" void rgmpm_rebuild_mapping_##SUF(void* h, int sort) {                                       \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.RebuildMapping(&r->state, sort != 0);                                                      \
  }                                                                                                      \
  extern "C" void rgmpm_calc_fem_##SUF(void* h, T dt) {                                                  \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.CalcFemStateAndForce(&r->state, dt);                                                       \
  }                                                                                                      \
  extern "C" void rgmpm_p2g_##SUF(void* h, T dt) {                                                       \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.ParticleToGrid(&r->state, dt);                                                             \
  }                                                                                                      \
  extern "C" void rgmpm_update_grid_##SUF(void* h, int bc) {                                             \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.UpdateGrid(&r->state, bc);                                                                 \
  }                                                                                                      \
  extern "C" void rgmpm_g2p_##SUF(void* h, T dt) {                                                       \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.GridToParticle(&r->state, dt);                                                             \
  }                                                                                                      \
  extern "C" void rgmpm_sync_##SUF(void* h) { static_cast<RefHandle<T>*>(h)->solver.GpuSync(); }         \
  extern "C" void rgmpm_set_contacts_##SUF(void* h, size_t n, const uint32_t* mpm_id,                    \
                                           const uint32_t* rigid_id, const T* dist, const T* normal,     \
                                           const T* pos, const T* rigid_v, const T* rigid_p) {           \
    ref_set_contacts<T>(static_cast<RefHandle<T>*>(h), n, mpm_id, rigid_id, dist, normal, pos, rigid_v,  \
                        rigid_p);                                                                        \
  }                                                                                                      \
  extern "C" void rgmpm_update_contact_##SUF(void* h, T dt, T mu, T k, T d, int exact) {                 \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->solver.UpdateContact(&r->state, 0, 0, dt, mu, k, d, false, exact != 0);                           \
  }                                                                                                      \
  extern "C" void rgmpm_set_num_bodies_##SUF(void* h, size_t n) {                                        \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->state.ReallocateExternelBodies(n);                                                                \
    r->state.external_forces_host().resize(n);                                                           \
  }                                                                                                      \
  extern "C" void rgmpm_body_impulses_##SUF(void* h, T* tau, T* f) {                                     \
    auto* r = static_cast<RefHandle<T>*>(h);                                                             \
    r->state.ExternelBodyForceToHost();                                                                  \
    const size_t n = r->state.num_external_bodies();                                                     \
    memcpy(tau, r->state.external_forces_host().F_Bq_W_tau.data(), n * 3 * sizeof(T));                   \
    memcpy(f, r->state.external_forces_host().F_Bq_W_f.data(), n * 3 * sizeof(T));                       \
  }                                                                                                      \
  extern "C" uint32_t rgmpm_touched_count_##SUF(void* h) {                                               \
    return static_cast<RefHandle<T>*>(h)->state.grid_touched_cnt_host();                                 \
  }                                                                                                      \
  extern "C" size_t rgmpm_n_particles_##SUF(void* h) {                                                   \
    return static_cast<RefHandle<T>*>(h)->state.n_particles();                                           \
  }                                                                                                      \
  extern "C" size_t rgmpm_read_##SUF(void* h, const char* name, void* out, size_t cap) {                 \
    return ref_read<T>(static_cast<RefHandle<T>*>(h), name, out, cap);                                   \
  }

RGMPM_DEFINE(f32, float)
RGMPM_DEFINE(f64, double)

extern "C" const char* rgmpm_build_flags() {
#ifdef RGMPM_FLAGS
  return RGMPM_FLAGS;
#else
  return "unknown";
#endif
}
