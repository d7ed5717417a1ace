// migrate.cuh -- particle migration between the states of a position-owned spatial decomposition
// (SURVEY 8e "after G2P particle migration: count, send/recv of packed records, compact"; VERDICT r01 row e2;
// extension: upstream is single-GPU, settings.h:40 G_DEVICE_COUNT = 1 is unused).
//
// Ownership of a VOLUMETRIC particle may change hands (cloth meshes stay whole: a triangle gathers its vertices by
// index, cuda_mpm_kernels.cuh:183-294).  A state is created with spare INACTIVE particles -- rest volume 0, parked
// at scattered points of a parking box: they carry no mass and no stress, so every transfer sees them as zeros -- and migration swaps
// roles: a leaver's record (pos 3, vel 3, C 9, volume, F 9, material index = 26 scalars) is copied out and its slot
// becomes inactive; an arrival takes over an inactive slot (and its local particle id).  The transport between the
// ranks is the host layer's (drake_b200/slab.py: torch.distributed).  Results do not depend on who owns a particle
// (shared grid blocks are summed by the exchange of halo.cuh), only the locality of the shared-block set does, so
// migration may lag: it runs every few substeps, between GridToParticle and the next RebuildMapping.
#pragma once
#include "state.cuh"

namespace gmpm {

constexpr int kMigrateRecord = 26;

struct ParkBox {
  double lo[3], hi[3];
};
// Where inactive particle q waits: a low-discrepancy point (R3 sequence) of the parking box, so that the massless
// particles never pile up in one grid block (a block with 10^5 particles is 10^3 P2G tiles to sum).
__host__ __device__ inline void park_position(const ParkBox& b, uint32_t q, double (&x)[3]) {
  const double a[3] = {0.8191725133961645, 0.6710436067037893, 0.5497004779019703};
  for (int d = 0; d < 3; ++d) {
    const double t = (double)(q + 1u) * a[d];
    x[d] = b.lo[d] + (t - floor(t)) * (b.hi[d] - b.lo[d]);
  }
}

template <typename T>
__global__ void migrate_out_kernel(uint32_t N, uint32_t n_cloth, int axis, T lo, T hi, ParkBox park,
                                   const int* __restrict__ pid, T* __restrict__ pos, T* __restrict__ vel,
                                   T* __restrict__ C, T* __restrict__ vol, T* __restrict__ Fv,
                                   const uint8_t* __restrict__ mat_id, uint32_t cap, uint32_t* __restrict__ count,
                                   T* __restrict__ records, int* __restrict__ out_pid) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= N) return;
  const uint32_t p = (uint32_t)pid[s];
  if (p < n_cloth) return;
  const T v0 = vol[s];
  if (!(v0 > T(0))) return;
  const T x = pos[(size_t)axis * N + s];
  if (x >= lo && x < hi) return;
  const uint32_t o = atomicAdd(count, 1u);
  if (o >= cap) return;  // no room in the caller's buffer: stays for the next call (count tells)
  const uint32_t q = p - n_cloth;
  T* r = records + (size_t)o * kMigrateRecord;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    r[d] = pos[(size_t)d * N + s];
    r[3 + d] = vel[(size_t)d * N + s];
  }
#pragma unroll
  for (int e = 0; e < 9; ++e) {
    r[6 + e] = C[(size_t)e * N + s];
    r[16 + e] = Fv[(size_t)q * 9 + e];
  }
  r[15] = v0;
  r[25] = T(mat_id[q]);
  out_pid[o] = (int)p;
  // the slot becomes an inactive particle
  double pk[3];
  park_position(park, q, pk);
  pos[s] = T(pk[0]); pos[(size_t)N + s] = T(pk[1]); pos[(size_t)2 * N + s] = T(pk[2]);
#pragma unroll
  for (int d = 0; d < 3; ++d) vel[(size_t)d * N + s] = T(0);
#pragma unroll
  for (int e = 0; e < 9; ++e) {
    C[(size_t)e * N + s] = T(0);
    Fv[(size_t)q * 9 + e] = (e % 4 == 0) ? T(1) : T(0);
  }
  vol[s] = T(0);
}

// slots of inactive volumetric particles (any order; the host sorts the list so that the assignment is reproducible)
template <typename T>
__global__ void migrate_free_slots_kernel(uint32_t N, uint32_t n_cloth, const int* __restrict__ pid,
                                          const T* __restrict__ vol, uint32_t* __restrict__ count,
                                          uint32_t* __restrict__ slots) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= N) return;
  if ((uint32_t)pid[s] < n_cloth || vol[s] > T(0)) return;
  slots[atomicAdd(count, 1u)] = s;
}

template <typename T>
__global__ void migrate_in_kernel(uint32_t n, uint32_t N, uint32_t n_cloth, const uint32_t* __restrict__ slots,
                                  const T* __restrict__ records, const int* __restrict__ pid, T* __restrict__ pos,
                                  T* __restrict__ vel, T* __restrict__ C, T* __restrict__ vol, T* __restrict__ Fv,
                                  uint8_t* __restrict__ mat_id, int* __restrict__ out_pid) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = slots[i];
  const uint32_t p = (uint32_t)pid[s], q = p - n_cloth;
  const T* r = records + (size_t)i * kMigrateRecord;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    pos[(size_t)d * N + s] = r[d];
    vel[(size_t)d * N + s] = r[3 + d];
  }
#pragma unroll
  for (int e = 0; e < 9; ++e) {
    C[(size_t)e * N + s] = r[6 + e];
    Fv[(size_t)q * 9 + e] = r[16 + e];
  }
  vol[s] = r[15];
  mat_id[q] = (uint8_t)r[25];
  out_pid[i] = (int)p;
}

// Dry run of a migration (changes nothing): how many active volumetric particles lie outside [lo, hi), by the slab
// that would receive them (destination = number of edges <= x, the owner_of() of slab.py), and how many inactive
// slots this state has left.  out[0 .. n_edges] = leavers per destination, out[n_edges + 1] = free slots.
constexpr int kMigrateMaxEdges = 63;
struct MigrateEdges {
  double e[kMigrateMaxEdges];
};
template <typename T>
__global__ void migrate_count_kernel(uint32_t N, uint32_t n_cloth, int axis, T lo, T hi, MigrateEdges edges, int n_edges,
                                     const int* __restrict__ pid, const T* __restrict__ pos, const T* __restrict__ vol,
                                     uint32_t* __restrict__ out) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  bool is_free = false;
  if (s < N && (uint32_t)pid[s] >= n_cloth) {
    if (!(vol[s] > T(0))) {
      is_free = true;
    } else {
      const T x = pos[(size_t)axis * N + s];
      if (!(x >= lo && x < hi)) {
        int d = 0;
        while (d < n_edges && edges.e[d] <= (double)x) ++d;
        atomicAdd(out + d, 1u);  // leavers are few
      }
    }
  }
  const int n_free = __syncthreads_count(is_free);  // spares are many: one atomic per CTA
  if (threadIdx.x == 0 && n_free) atomicAdd(out + n_edges + 1, (uint32_t)n_free);
}

template <typename T>
void State<T>::migrate_count(int axis, T lo, T hi, const double* edges, size_t n_edges, size_t* leavers, size_t* n_free) {
  require_finalized();
  if (axis < 0 || axis > 2) throw std::invalid_argument("migrate_count: axis must be 0, 1 or 2");
  if (n_edges > (size_t)kMigrateMaxEdges) throw std::invalid_argument("migrate_count: at most 63 slab edges");
  if ((n_edges && !edges) || !leavers || !n_free) throw std::invalid_argument("migrate_count: null argument");
  for (size_t d = 0; d <= n_edges; ++d) leavers[d] = 0;
  *n_free = 0;
  if (!n_vol) return;
  MigrateEdges e{};
  for (size_t d = 0; d < n_edges; ++d) {
    e.e[d] = edges[d];
    if (d && !(e.e[d] >= e.e[d - 1])) throw std::invalid_argument("migrate_count: edges must ascend");
  }
  uint32_t* d_out = static_cast<uint32_t*>(io(sizeof(uint32_t) * (kMigrateMaxEdges + 2)));
  GMPM_CUDA(cudaMemsetAsync(d_out, 0, sizeof(uint32_t) * (n_edges + 2), stream));
  launch("migrate_count");
  migrate_count_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, n_cloth(), axis, lo, hi, e, (int)n_edges,
                                                                 pid[cur], pos[cur], vol[cur], d_out);
  GMPM_CUDA(cudaGetLastError());
  uint32_t h[kMigrateMaxEdges + 2];
  to_host(h, d_out, sizeof(uint32_t) * (n_edges + 2));
  for (size_t d = 0; d <= n_edges; ++d) leavers[d] = h[d];
  *n_free = h[n_edges + 1];
}

template <typename T>
void State<T>::invalidate_after_particle_edit() {
  keys_valid = bin_valid = fem_valid = vis_keys_valid = false;
  snapshot_valid = false;
  n_contacts = 0;
  cb_valid = false;
  g2p_queue_ready = false;
  halo_planned = halo_plan_pending = halo_flags_out = halo_own_valid = halo_packed_out = false;
  halo_sent = halo_recv_pending = false;
  this->stage = GridStage::kClean;
}

template <typename T>
size_t State<T>::migrate_out(int axis, T lo, T hi, const T* park, T* records, int* pids_out, size_t cap) {
  require_finalized();
  if (axis < 0 || axis > 2) throw std::invalid_argument("migrate_out: axis must be 0, 1 or 2");
  if (!park || (cap && (!records || !pids_out))) throw std::invalid_argument("migrate_out: null argument");
  if (!n_vol || !cap) return 0;
  const size_t rec_bytes = sizeof(T) * kMigrateRecord * cap;
  char* buf = static_cast<char*>(io(rec_bytes + sizeof(int) * cap + 16));
  uint32_t* d_count = reinterpret_cast<uint32_t*>(buf);
  T* d_rec = reinterpret_cast<T*>(buf + 16);
  int* d_pid = reinterpret_cast<int*>(buf + 16 + rec_bytes);
  GMPM_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), stream));
  launch("migrate_out");
  ParkBox box;
  for (int d = 0; d < 3; ++d) {
    box.lo[d] = (double)park[d];
    box.hi[d] = (double)park[3 + d];
  }
  migrate_out_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, n_cloth(), axis, lo, hi, box,
                                                               pid[cur], pos[cur], vel[cur], C[cur], vol[cur], Fv, mat_id,
                                                               (uint32_t)cap, d_count, d_rec, d_pid);
  GMPM_CUDA(cudaGetLastError());
  uint32_t n = 0;
  to_host(&n, d_count, sizeof(uint32_t));
  n = (uint32_t)std::min<size_t>(n, cap);
  if (n) {
    // atomics hand out the rows in any order: sort by local particle id so that a run is reproducible
    std::vector<T> r((size_t)n * kMigrateRecord);
    std::vector<int> p(n);
    to_host(r.data(), d_rec, sizeof(T) * r.size());
    to_host(p.data(), d_pid, sizeof(int) * n);
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return p[a] < p[b]; });
    for (uint32_t i = 0; i < n; ++i) {
      pids_out[i] = p[order[i]];
      std::memcpy(records + (size_t)i * kMigrateRecord, r.data() + (size_t)order[i] * kMigrateRecord, sizeof(T) * kMigrateRecord);
    }
    invalidate_after_particle_edit();
  }
  return n;
}

template <typename T>
void State<T>::migrate_in(const T* records, size_t n, int* pids_out) {
  require_finalized();
  if (!n) return;
  if (!records) throw std::invalid_argument("migrate_in: null records");
  if (!n_vol) throw std::invalid_argument("migrate_in: the state holds no volumetric particles (no spare slots)");
  // the material index of a record is an index into THIS state's table (it is rank-local: all states of a scene
  // must register their materials in the same order); an index outside the table would read past the kernel's
  // by-value material array
  for (size_t i = 0; i < n; ++i) {
    const double m = (double)records[i * kMigrateRecord + 25];
    if (!(m >= 0.0) || m >= (double)materials.size() || m != std::floor(m))
      throw std::invalid_argument("migrate_in: record " + std::to_string(i) + " carries material index " +
                                  std::to_string(m) + " but this state has " + std::to_string(materials.size()) +
                                  " material(s) registered (all states of a scene must add their materials in the same order)");
  }
  char* buf = static_cast<char*>(io(16 + sizeof(uint32_t) * N + sizeof(T) * kMigrateRecord * n + sizeof(int) * n));
  uint32_t* d_count = reinterpret_cast<uint32_t*>(buf);
  uint32_t* d_slots = reinterpret_cast<uint32_t*>(buf + 16);
  T* d_rec = reinterpret_cast<T*>(buf + 16 + ((sizeof(uint32_t) * N + 15) & ~(size_t)15));
  int* d_pid = reinterpret_cast<int*>(reinterpret_cast<char*>(d_rec) + sizeof(T) * kMigrateRecord * n);
  GMPM_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), stream));
  launch("migrate_free_slots");
  migrate_free_slots_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, n_cloth(), pid[cur], vol[cur], d_count, d_slots);
  GMPM_CUDA(cudaGetLastError());
  uint32_t n_free = 0;
  to_host(&n_free, d_count, sizeof(uint32_t));
  if (n_free < n)
    throw std::runtime_error("migrate_in: " + std::to_string(n) + " arrivals but only " + std::to_string(n_free) +
                             " inactive particles left: create the state with more spare particles (volume 0)");
  std::vector<uint32_t> slots(n_free);
  to_host(slots.data(), d_slots, sizeof(uint32_t) * n_free);
  // the n lowest free slots, ascending (reproducible assignment) -- without sorting the whole free list
  std::nth_element(slots.begin(), slots.begin() + (n - 1), slots.end());
  std::sort(slots.begin(), slots.begin() + n);
  GMPM_CUDA(cudaMemcpyAsync(d_slots, slots.data(), sizeof(uint32_t) * n, cudaMemcpyHostToDevice, stream));
  GMPM_CUDA(cudaMemcpyAsync(d_rec, records, sizeof(T) * kMigrateRecord * n, cudaMemcpyHostToDevice, stream));
  launch("migrate_in");
  migrate_in_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>((uint32_t)n, (uint32_t)N, n_cloth(), d_slots, d_rec, pid[cur], pos[cur],
                                                              vel[cur], C[cur], vol[cur], Fv, mat_id, d_pid);
  GMPM_CUDA(cudaGetLastError());
  if (pids_out) to_host(pids_out, d_pid, sizeof(int) * n);
  else GMPM_CUDA(cudaStreamSynchronize(stream));  // slots / records are borrowed host memory
  invalidate_after_particle_edit();
}

}  // namespace gmpm
