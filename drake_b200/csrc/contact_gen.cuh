// contact_gen.cuh -- device-side generation of the MPM-rigid contact pairs for ANALYTIC rigid geometry.
//
// In the reference the pairs are built on the host every substep: DeformableDriver::CalcMpmContactPairs
// (multibody/plant/deformable_driver.h:120-194) downloads all particle positions, asks
// GeometryState::ComputeSignedDistanceToPoint for every particle and uploads seven arrays
// (GpuMpmSolver::CopyContactPairs, cuda_mpm_solver.cu:194-212).  For boxes and spheres the signed distance
// is a closed form, so the same pairs can be produced where the positions already are.
// The result is what the host path would upload, in the same order (geometry by geometry, particles in
// visible order), from the same position snapshot (the last SyncParticleStateToCpu), so UpdateContact sees
// identical input.  General meshes / FCL shapes stay on the host path.
#pragma once
#include "contact.cuh"

namespace gmpm {

constexpr int kGenThreads = 256;
constexpr int kGenItemsPerThread = 4;
constexpr int kGenItemsPerCta = kGenThreads * kGenItemsPerThread;

template <typename T>
struct GeomDev {
  int kind;           // 0 box [a, b], 1 sphere (centre a, radius b[0])
  uint32_t rigid_id;
  T a[3], b[3], v[3], p[3];
};

// Signed distance query of one point (what ComputeSignedDistanceToPoint returns for these shapes):
// inside -> true, dist < 0, grad = outward unit normal of the nearest surface point.
template <typename T>
__device__ __forceinline__ bool analytic_inside(const GeomDev<T>& g, const T* x, T& dist, T* grad) {
  grad[0] = grad[1] = grad[2] = T(0);
  if (g.kind == 0) {
    if (!(x[0] > g.a[0] && x[0] < g.b[0] && x[1] > g.a[1] && x[1] < g.b[1] && x[2] > g.a[2] && x[2] < g.b[2]))
      return false;
    T best = x[0] - g.a[0];
    int face = 0;
#pragma unroll
    for (int f = 1; f < 6; ++f) {  // first minimum in the order x-lo, y-lo, z-lo, hi-x, hi-y, hi-z
      const T d = f < 3 ? x[f] - g.a[f] : g.b[f - 3] - x[f - 3];
      if (d < best) {
        best = d;
        face = f;
      }
    }
    dist = -best;
    grad[face % 3] = face < 3 ? T(-1) : T(1);
    return true;
  }
  const T r[3] = {x[0] - g.a[0], x[1] - g.a[1], x[2] - g.a[2]};
  const T len = tsqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (!(len < g.b[0])) return false;
  dist = len - g.b[0];
  if (len > T(0)) {
    grad[0] = r[0] / len;
    grad[1] = r[1] / len;
    grad[2] = r[2] / len;
  } else {
    grad[2] = T(1);
  }
  return true;
}

// Item index = geometry * N + visible particle.  Pass 1 counts the hits of each CTA's span; pass 2 repeats
// the test and writes the pairs at their rank (stable: CTA offset = sum of the preceding counts).
template <typename T, bool WRITE>
__global__ void __launch_bounds__(kGenThreads)
contact_gen_kernel(uint32_t N, uint32_t n_geoms, const GeomDev<T>* __restrict__ geoms, const T* __restrict__ pos_vis,
                   uint32_t* __restrict__ cta_counts, uint32_t* __restrict__ total, uint32_t* __restrict__ c_pid,
                   uint32_t* __restrict__ c_rigid_id, T* __restrict__ c_dist, T* __restrict__ c_normal,
                   T* __restrict__ c_pos, T* __restrict__ c_rigid_v, T* __restrict__ c_rigid_p) {
  __shared__ uint32_t s_warp[kGenThreads / 32];
  __shared__ uint32_t s_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t n_items = (uint64_t)N * n_geoms;
  const uint64_t i0 = ((uint64_t)blockIdx.x * kGenThreads + threadIdx.x) * kGenItemsPerThread;
  bool hit[kGenItemsPerThread];
  T dist[kGenItemsPerThread], grad[kGenItemsPerThread][3], x[kGenItemsPerThread][3];
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < kGenItemsPerThread; ++j) {
    hit[j] = false;
    const uint64_t i = i0 + j;
    if (i < n_items) {
      const uint32_t g = (uint32_t)(i / N), v = (uint32_t)(i % N);
      x[j][0] = pos_vis[(size_t)v * 3];
      x[j][1] = pos_vis[(size_t)v * 3 + 1];
      x[j][2] = pos_vis[(size_t)v * 3 + 2];
      hit[j] = analytic_inside(geoms[g], x[j], dist[j], grad[j]);
    }
    c += hit[j] ? 1u : 0u;
  }
  if (!WRITE) {
    c = __reduce_add_sync(kFullMask, c);
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    if (lane == 0 && c) atomicAdd(&s_base, c);
    __syncthreads();
    if (threadIdx.x == 0) cta_counts[blockIdx.x] = s_base;
    return;
  }
  uint32_t pre = 0;
  for (uint32_t b = threadIdx.x; b < blockIdx.x; b += kGenThreads) pre += cta_counts[b];
  pre = __reduce_add_sync(kFullMask, pre);
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  if (lane == 0 && pre) atomicAdd(&s_base, pre);
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(kFullMask, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t off = s_base + incl - c;
  for (int w = 0; w < warp; ++w) off += s_warp[w];
#pragma unroll
  for (int j = 0; j < kGenItemsPerThread; ++j) {
    if (!hit[j]) continue;
    const uint64_t i = i0 + j;
    const GeomDev<T>& g = geoms[(uint32_t)(i / N)];
    c_pid[off] = (uint32_t)(i % N);
    c_rigid_id[off] = g.rigid_id;
    c_dist[off] = dist[j];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      c_normal[(size_t)off * 3 + d] = -grad[j][d];  // deformable_driver.h:184: normal = -grad_W
      c_pos[(size_t)off * 3 + d] = x[j][d];
      c_rigid_v[(size_t)off * 3 + d] = g.v[d];
      c_rigid_p[(size_t)off * 3 + d] = g.p[d];
    }
    ++off;
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kGenThreads - 1) *total = off;
}

template <typename T>
size_t State<T>::generate_contact_pairs(const gmpm_analytic_geometry* geoms, size_t n_geoms) {
  require_finalized();
  cb_valid = false;
  n_contacts = 0;
  if (n_geoms == 0 || N == 0) return 0;
  if (!geoms) throw std::invalid_argument("generate_contact_pairs: null geometry array");
  if (n_geoms > 64) throw std::invalid_argument("generate_contact_pairs: at most 64 analytic geometries");
  std::vector<GeomDev<T>> h(n_geoms);
  for (size_t i = 0; i < n_geoms; ++i) {
    const gmpm_analytic_geometry& g = geoms[i];
    if (g.kind != GMPM_GEOM_BOX && g.kind != GMPM_GEOM_SPHERE) throw std::invalid_argument("unknown analytic geometry kind");
    h[i].kind = g.kind;
    h[i].rigid_id = g.rigid_id;
    for (int d = 0; d < 3; ++d) {
      h[i].a[d] = T(g.lo[d]);
      h[i].b[d] = T(g.hi[d]);
      h[i].v[d] = T(g.velocity[d]);
      h[i].p[d] = T(g.p_WB[d]);
    }
  }
  if (!snapshot_valid) snapshot_positions_device();  // same data the host query would read: the last position snapshot
  const uint64_t n_items = (uint64_t)N * n_geoms;
  const uint32_t ctas = (uint32_t)((n_items + kGenItemsPerCta - 1) / kGenItemsPerCta);
  if (ctas > gen_cap) {
    GMPM_CUDA(cudaStreamSynchronize(stream));
    if (gen_counts) dfree(gen_counts);
    if (gen_geoms) dfree(gen_geoms);
    gen_counts = dalloc<uint32_t>(ctas + 1, false);
    gen_geoms = dalloc<unsigned char>(64 * sizeof(GeomDev<double>), false);
    gen_cap = ctas;
  }
  GeomDev<T>* dg = reinterpret_cast<GeomDev<T>*>(gen_geoms);
  GMPM_CUDA(cudaMemcpyAsync(dg, h.data(), sizeof(GeomDev<T>) * n_geoms, cudaMemcpyHostToDevice, stream));
  GMPM_CUDA(cudaStreamSynchronize(stream));  // h goes out of scope; also orders the snapshot gather
  uint32_t* total = gen_counts + ctas;
  launch("contact_gen_count");
  contact_gen_kernel<T, false><<<ctas, kGenThreads, 0, stream>>>((uint32_t)N, (uint32_t)n_geoms, dg, pos_stage, gen_counts,
                                                                total, nullptr, nullptr, nullptr, nullptr, nullptr,
                                                                nullptr, nullptr);
  GMPM_CUDA(cudaGetLastError());
  // the number of pairs sizes the (grow-only) contact buffers: one small read-back, as ReallocateContacts needs it
  std::vector<uint32_t> hc(ctas);
  to_host(hc.data(), gen_counts, sizeof(uint32_t) * ctas);
  size_t n = 0;
  for (uint32_t c : hc) n += c;
  if (n == 0) return 0;
  if (n >= (size_t)1 << 31) throw std::invalid_argument("too many contact pairs");
  reallocate_contacts(n);
  launch("contact_gen_write");
  contact_gen_kernel<T, true><<<ctas, kGenThreads, 0, stream>>>((uint32_t)N, (uint32_t)n_geoms, dg, pos_stage, gen_counts,
                                                               total, c_pid, c_rigid_id, c_dist, c_normal, c_pos,
                                                               c_rigid_v, c_rigid_p);
  n_contacts = n;
  launch("contact_init_vel");
  contact_init_vel_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>((uint32_t)n, (uint32_t)N, c_pid,
                                                                    pid_of_vis[vis_cur], slot_of_pid, vel[cur], c_vel);
  GMPM_CUDA(cudaGetLastError());
  return n;
}

// The DeformableDriver substep loop (deformable_driver.h:240-260) on the library side of the ABI, for scenes whose
// rigid geometry is analytic: snapshot -> RebuildMapping -> FEM -> P2G -> UpdateGrid -> contact query ->
// UpdateContact -> G2P, n times in one call.  `query_every` is MpmConfigParams::contact_query_frequency
// (cpu_mpm_model.h:23, declared and never read upstream): the pairs are regenerated every query_every-th substep
// (and after every visible re-sort, which renumbers the particles) and reused in between with their contact
// velocities re-initialised.  The geometries move with their `velocity` from the pose they were given.  The
// position snapshot is taken AFTER RebuildMapping: pair indices are visible indices and a re-sort renumbers
// them (upstream's driver never sorts in this loop, deformable_driver.h:244).
template <typename T>
void State<T>::run_driver_substeps(int n, T dt, int bc, int sort_every, const gmpm_analytic_geometry* geoms,
                                   size_t n_geoms, int query_every, T mu, T stiffness, T damping, bool exact,
                                   int max_it, gmpm_driver_stats* out) {
  require_finalized();
  if (n_geoms && !geoms) throw std::invalid_argument("run_driver_substeps: null geometry array");
  if (query_every < 1) query_every = 1;
  gmpm_driver_stats st{};
  std::vector<gmpm_analytic_geometry> moved(geoms, geoms + n_geoms);
  bool have_pairs = false;
  for (int i = 0; i < n; ++i) {
    const bool sort = sort_every > 0 && (i % sort_every) == 0;
    const bool query = n_geoms && (!have_pairs || sort || (i % query_every) == 0);
    rebuild_mapping(sort);
    if (query) snapshot_positions_device();  // the positions SyncParticleStateToCpu would hand to the host query (:243)
    calc_fem(dt);
    p2g(dt);
    update_grid(bc);
    if (query) {
      for (size_t g = 0; g < n_geoms; ++g)
        for (int d = 0; d < 3; ++d) {
          const double shift = driver_geometry_frozen ? 0.0 : geoms[g].velocity[d] * (double)dt * i;
          moved[g].lo[d] = geoms[g].lo[d] + shift;
          if (geoms[g].kind == GMPM_GEOM_BOX) moved[g].hi[d] = geoms[g].hi[d] + shift;
          moved[g].p_WB[d] = geoms[g].p_WB[d] + shift;
        }
      generate_contact_pairs(moved.data(), n_geoms);
      have_pairs = true;
      ++st.contact_queries;
    } else if (n_contacts) {
      launch("contact_init_vel");
      contact_init_vel_kernel<T><<<blocks_for(n_contacts), kThreads, 0, stream>>>(
          (uint32_t)n_contacts, (uint32_t)N, c_pid, pid_of_vis[vis_cur], slot_of_pid, vel[cur], c_vel);
      GMPM_CUDA(cudaGetLastError());
    }
    if (n_contacts) {
      gmpm_contact_stats cs{};
      update_contact(dt, mu, stiffness, damping, exact, max_it, &cs);
      st.newton_iterations += cs.iterations;
      if (cs.iterations > st.newton_iterations_max) st.newton_iterations_max = cs.iterations;
      st.last_residual = cs.residual;
    }
    st.contact_pairs += n_contacts;
    g2p(dt);
    ++st.substeps;
  }
  if (out) *out = st;
}

}  // namespace gmpm
