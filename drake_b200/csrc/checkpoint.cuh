// checkpoint.cuh -- full checkpoint / restore of the particle and face state.
//
// The reference cannot resume an MPM run: its state is raw device pointers inside an AbstractValue, and
// DumpCpuState returns vertex positions and indices only (cuda_mpm_model.cu:245-265) -- velocities, the
// affine matrices and the deformation gradients are lost.  A checkpoint here is one host blob in the
// REFERENCE's layout and visible particle order (so it does not depend on the physical order the library
// happens to keep): header, positions[3N] velocities[3N] volumes[N] affine[9N] F[9F] Dm_inv[4F] particle_F[9Nv] pids[N]
// material_index[Nv] (bytes, by particle id like particle_F, padded to 8: a migration rewrites the material index of
// the slot an arrival takes, so it is state, not something the AddParticles calls of the rebuilt scene reproduce).
// Loading needs a finalized state built from the same meshes (same N, F, scalar); the run then continues
// with the same arithmetic on a different physical order, i.e. equal to summation-order rounding.
#pragma once
#include "getters.cuh"

namespace gmpm {

struct CheckpointHeader {
  char magic[8];          // "GMPMCKP3" (2: without the material indices)
  uint32_t scalar_bytes;  // 4 or 8
  uint32_t domain_bits;
  uint64_t N, n_verts, n_faces, n_vol;
  double dx;
};

// reference layout, visible order -> planar by slot, with slot = visible index
template <typename T>
__global__ void scatter_particle_field_kernel(uint32_t N, int comps, const T* __restrict__ in, T* __restrict__ out) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  for (int c = 0; c < comps; ++c) out[(size_t)c * N + v] = in[(size_t)v * comps + c];
}
__global__ void restore_order_kernel(uint32_t N, const int* __restrict__ pids, int* __restrict__ pid_slot,
                                     int* __restrict__ slot_of_pid, int* __restrict__ pid_of_vis,
                                     int* __restrict__ vis_of_pid) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const int p = pids[v];
  pid_slot[v] = p;
  slot_of_pid[p] = (int)v;
  pid_of_vis[v] = p;
  vis_of_pid[p] = (int)v;
}
__global__ void iota_kernel(uint32_t n, int* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int)i;
}

template <typename T>
size_t State<T>::checkpoint_bytes() const {
  return sizeof(CheckpointHeader) + sizeof(T) * (16 * N + 13 * n_faces + 9 * n_vol) + sizeof(int) * N +
         ((n_vol + 7) & ~(size_t)7);
}

template <typename T>
size_t State<T>::save_state(void* out, size_t cap) {
  require_finalized();
  const size_t total = checkpoint_bytes();
  if (!out || cap < total) throw std::invalid_argument("gmpm_save_state: buffer too small (see gmpm_checkpoint_bytes)");
  CheckpointHeader h{};
  std::memcpy(h.magic, "GMPMCKP3", 8);
  h.scalar_bytes = sizeof(T);
  h.domain_bits = (uint32_t)k.domain_bits;
  h.N = N;
  h.n_verts = n_verts;
  h.n_faces = n_faces;
  h.n_vol = n_vol;
  h.dx = (double)k.dx;
  char* p = static_cast<char*>(out);
  std::memcpy(p, &h, sizeof(h));
  p += sizeof(h);
  for (const char* name : {"positions", "velocities", "volumes", "affine_matrices", "deformation_gradients",
                           "Dm_inverses", "particle_F", "pids"})
    p += read(name, p, cap - (size_t)(p - static_cast<char*>(out)));
  {
    const size_t padded = (n_vol + 7) & ~(size_t)7;
    std::memset(p, 0, padded);
    if (n_vol) to_host(p, mat_id, n_vol);
    p += padded;
  }
  if ((size_t)(p - static_cast<char*>(out)) != total) throw std::runtime_error("gmpm_save_state: size mismatch");
  return total;
}

template <typename T>
void State<T>::load_state(const void* in, size_t bytes) {
  require_finalized();
  if (!in || bytes < sizeof(CheckpointHeader)) throw std::invalid_argument("gmpm_load_state: not a checkpoint");
  CheckpointHeader h;
  std::memcpy(&h, in, sizeof(h));
  if (std::memcmp(h.magic, "GMPMCKP2", 8) == 0)
    throw std::invalid_argument("gmpm_load_state: version-2 checkpoint (no material indices): save it again with this library");
  if (std::memcmp(h.magic, "GMPMCKP3", 8) != 0) throw std::invalid_argument("gmpm_load_state: bad magic");
  if (h.scalar_bytes != sizeof(T) || h.N != N || h.n_verts != n_verts || h.n_faces != n_faces || h.n_vol != n_vol ||
      h.domain_bits != (uint32_t)k.domain_bits || bytes != checkpoint_bytes())
    throw std::invalid_argument("gmpm_load_state: the checkpoint was taken from a different scene (sizes, scalar or grid)");
  if (std::fabs(h.dx - (double)k.dx) > 1e-12 * std::fabs((double)k.dx))
    throw std::invalid_argument("gmpm_load_state: the checkpoint was taken with a different cell size dx");
  if (N == 0) return;
  const uint32_t n = (uint32_t)N, F = (uint32_t)n_faces;
  const char* p = static_cast<const char*>(in) + sizeof(h);
  {
    // pids drive scattered device writes: they must be a permutation of 0..N-1 that keeps face particles (pid <
    // n_faces) distinguishable -- checked on the host, where the blob already is
    const int* hp = reinterpret_cast<const int*>(p + sizeof(T) * (16 * N + 13 * n_faces + 9 * n_vol));
    std::vector<uint8_t> seen(N, 0);
    for (size_t v = 0; v < N; ++v) {
      int q;
      std::memcpy(&q, hp + v, sizeof(int));
      if (q < 0 || (size_t)q >= N || seen[(size_t)q])
        throw std::invalid_argument("gmpm_load_state: corrupt checkpoint (pids are not a permutation of 0..N-1)");
      seen[(size_t)q] = 1;
    }
  }
  const uint8_t* h_mat = reinterpret_cast<const uint8_t*>(p + sizeof(T) * (16 * N + 13 * n_faces + 9 * n_vol) + sizeof(int) * N);
  for (size_t q = 0; q < n_vol; ++q)
    if (h_mat[q] >= materials.size())
      throw std::invalid_argument("gmpm_load_state: the checkpoint refers to material " + std::to_string((int)h_mat[q]) +
                                  " but this state has " + std::to_string(materials.size()) +
                                  " registered (rebuild the scene with the same AddParticles materials, in the same order)");
  // one staging upload, then layout conversion on the device (slot = visible index)
  const size_t payload = bytes - sizeof(h);
  char* stage = static_cast<char*>(io(payload));
  GMPM_CUDA(cudaMemcpyAsync(stage, p, payload, cudaMemcpyHostToDevice, stream));
  const T* s_pos = reinterpret_cast<const T*>(stage);
  const T* s_vel = s_pos + 3 * N;
  const T* s_vol = s_vel + 3 * N;
  const T* s_C = s_vol + N;
  const T* s_F = s_C + 9 * N;
  const T* s_Dm = s_F + 9 * n_faces;
  const T* s_Fv = s_Dm + 4 * n_faces;
  const int* s_pid = reinterpret_cast<const int*>(s_Fv + 9 * n_vol);  // ints last: every T array stays aligned
  launch("restore_particles");
  scatter_particle_field_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>(n, 3, s_pos, pos[cur]);
  scatter_particle_field_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>(n, 3, s_vel, vel[cur]);
  scatter_particle_field_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>(n, 1, s_vol, vol[cur]);
  scatter_particle_field_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>(n, 9, s_C, C[cur]);
  restore_order_kernel<<<blocks_for(n), kThreads, 0, stream>>>(n, s_pid, pid[cur], slot_of_pid, pid_of_vis[vis_cur],
                                                              vis_of_pid);
  if (F) {
    // face records back in face-id order (fid = identity); tri is unchanged by a run
    std::vector<int> ht(3 * (size_t)F);
    for (size_t f = 0; f < F; ++f)
      for (int c = 0; c < 3; ++c) ht[c * (size_t)F + f] = h_tri[f * 3 + c] + (int)F;
    GMPM_CUDA(cudaMemcpyAsync(tri, ht.data(), sizeof(int) * ht.size(), cudaMemcpyHostToDevice, stream));
    iota_kernel<<<blocks_for(F), kThreads, 0, stream>>>(F, fid);
    scatter_particle_field_kernel<T><<<blocks_for(F), kThreads, 0, stream>>>(F, 9, s_F, Fd);
    scatter_particle_field_kernel<T><<<blocks_for(F), kThreads, 0, stream>>>(F, 4, s_Dm, Dm_inv);
    launches += 3;
  }
  if (n_vol)  // deformation gradients of the volumetric particles: by particle id on both sides
  {
    GMPM_CUDA(cudaMemcpyAsync(Fv, s_Fv, sizeof(T) * 9 * n_vol, cudaMemcpyDeviceToDevice, stream));
    GMPM_CUDA(cudaMemcpyAsync(mat_id, reinterpret_cast<const uint8_t*>(s_pid + N), n_vol, cudaMemcpyDeviceToDevice, stream));
  }
  launches += 4;
  GMPM_CUDA(cudaGetLastError());
  GMPM_CUDA(cudaStreamSynchronize(stream));  // the host blob and ht are borrowed
  keys_valid = bin_valid = fem_valid = vis_keys_valid = false;
  forces_zero = false;
  snapshot_valid = false;
  fem_calls = 0;       // the face records are in face-id order again: re-sort them after the first G2P
  n_contacts = 0;
  cb_valid = false;
  g2p_queue_ready = false;
  halo_planned = halo_plan_pending = halo_flags_out = halo_own_valid = halo_packed_out = false;
  halo_sent = halo_recv_pending = false;
  this->stage = GridStage::kClean;   // whatever the grid held belongs to the replaced particles: the next P2G starts over
}

}  // namespace gmpm
