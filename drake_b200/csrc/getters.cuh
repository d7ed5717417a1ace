// getters.cuh -- State<T>::read: copies a named buffer to the host in the REFERENCE's layout and
// visible particle order (replaces the raw device-pointer accessors of cuda_mpm_model.cuh:43-143,
// which tests and the driver use through cudaMemcpy).
#pragma once
#include "state.cuh"

namespace gmpm {

template <typename T>
__global__ void mask_non_face_rows_kernel(uint32_t N, uint32_t n_faces, int comps, const int* __restrict__ pid_of_vis,
                                          T* __restrict__ out) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N || (uint32_t)pid_of_vis[v] < n_faces) return;
  for (int c = 0; c < comps; ++c) out[(size_t)v * comps + c] = T(0);
}

__global__ void contact_vis_keys_kernel(uint32_t n, const uint32_t* __restrict__ keys, uint32_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[i];
}

template <typename T>
size_t State<T>::read(const char* name, void* out, size_t cap) {
  require_finalized();
  const std::string n(name);
  auto need = [&](size_t bytes) {
    if (bytes > cap) throw std::invalid_argument("gmpm_read(" + n + "): buffer too small");
    return bytes;
  };
  auto particle_field = [&](const T* src, int comps, bool faces_only) -> size_t {
    const size_t bytes = need(sizeof(T) * comps * N);
    if (!N) return 0;
    T* buf = static_cast<T*>(io(bytes));
    launch("gather_particle_field");
    gather_particle_field_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, comps, src,
                                                                           pid_of_vis[vis_cur], slot_of_pid, buf);
    if (faces_only) {
      launch("mask_non_face_rows");
      mask_non_face_rows_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, (uint32_t)n_faces, comps,
                                                                          pid_of_vis[vis_cur], buf);
    }
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  };
  auto face_field = [&](const T* src, int comps) -> size_t {
    const size_t bytes = need(sizeof(T) * comps * n_faces);
    if (!n_faces) return 0;
    T* buf = static_cast<T*>(io(bytes));
    launch("transpose_planar");
    transpose_planar_kernel<T><<<blocks_for(n_faces), kThreads, 0, stream>>>((uint32_t)n_faces, comps, src, fid, buf);
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  };
  auto grid_vec = [&](const T* src) -> size_t {
    const size_t bytes = need(sizeof(T) * 3 * k.n_cells);
    T* buf = static_cast<T*>(io(bytes));
    launch("grid_vec_to_reference");
    grid_vec_to_reference_kernel<T><<<blocks_for(k.n_cells), kThreads, 0, stream>>>(k.n_cells, src, buf);
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  };
  auto plain = [&](const void* src, size_t bytes) -> size_t {
    need(bytes);
    if (bytes) to_host(out, src, bytes);
    return bytes;
  };

  if (n == "positions") return particle_field(pos[cur], 3, false);
  if (n == "velocities") return particle_field(vel[cur], 3, false);
  if (n == "volumes") return particle_field(vol[cur], 1, false);
  if (n == "affine_matrices") return particle_field(C[cur], 9, false);
  if ((n == "forces" || n == "taus") && !fem_valid)
    throw std::invalid_argument("forces / taus are defined between CalcFemStateAndForce and GridToParticle only "
                                "(the reference overwrites them in the next CalcFemStateAndForce)");
  if (n == "forces") return particle_field(forces, 3, false);
  if (n == "taus") {
    const size_t bytes = need(sizeof(T) * 9 * N);
    if (!N) return 0;
    T* buf = static_cast<T*>(io(bytes));
    launch("gather_taus");
    gather_taus_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, (uint32_t)n_faces, taus,
                                                                 pid_of_vis[vis_cur], slot_of_pid, buf, n_cloth());
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  }
  if (n == "particle_F") return plain(Fv, sizeof(T) * 9 * n_vol);  // volumetric particles, by particle id (extension)
  if (n == "dev_counters") return plain(ctr, sizeof(DevCounters));  // instrumentation (tools/)
  if (n == "pids") return plain(pid_of_vis[vis_cur], sizeof(int) * N);
  if (n == "index_mappings") return plain(vis_of_pid, sizeof(int) * N);
  if (n == "sort_keys") {
    materialize_visible_keys();
    return plain(vis_keys[vis_res], sizeof(uint32_t) * N);
  }
  if (n == "sort_ids") {
    materialize_visible_keys();
    return plain(vis_ids[vis_res], sizeof(uint32_t) * N);
  }
  if (n == "indices") {
    const size_t bytes = need(sizeof(int) * 3 * n_faces);
    if (!n_faces) return 0;
    int* buf = static_cast<int*>(io(bytes));
    launch("transpose_planar");
    transpose_planar_kernel<int><<<blocks_for(n_faces), kThreads, 0, stream>>>((uint32_t)n_faces, 3, tri, fid, buf);
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  }
  if (n == "Dm_inverses") return face_field(Dm_inv, 4);
  if (n == "deformation_gradients") return face_field(Fd, 9);
  if (n == "grid_masses") {
    assemble_momentum();
    return plain(g_mass, sizeof(T) * k.n_cells);
  }
  if (n == "grid_momentum") {
    assemble_momentum();
    return grid_vec(g_vel);
  }
  if (n == "grid_v_star") return grid_vec(g_vstar);
  if (n == "grid_touched_flags") {
    const size_t bytes = need(sizeof(uint32_t) * k.n_blocks);
    uint32_t* buf = static_cast<uint32_t*>(io(bytes));
    GMPM_CUDA(cudaMemsetAsync(buf, 0, bytes, stream));
    launch("flags_from_list");
    flags_from_list_kernel<<<blocks_for(k.n_blocks), kThreads, 0, stream>>>(touched_list[tl_cur], ctr, buf);
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, bytes);
    return bytes;
  }
  if (n == "grid_touched_ids") return plain(touched_list[tl_cur], sizeof(uint32_t) * k.n_blocks);
  if (n == "grid_Hess" || n == "grid_Grad" || n == "grid_Dir") {
    const int comps = n == "grid_Hess" ? 9 : 3;
    const T* src = n == "grid_Hess" ? g_Hess : (n == "grid_Grad" ? g_Grad : g_Dir);
    const size_t bytes = need(sizeof(T) * comps * k.n_cells);
    const T* buf = contact_grid_dense(src, comps);
    to_host(out, buf, bytes);
    return bytes;
  }
  if (n == "contact_vel") return plain(c_vel, sizeof(T) * 3 * n_contacts);
  if (n == "contact_vel0") return plain(c_vel0, sizeof(T) * 3 * n_contacts);
  if (n == "contact_sort_keys") return plain(c_keys, sizeof(uint32_t) * n_contacts);
  // the contact-pair arrays as uploaded by CopyContactPairs or produced by gmpm_generate_contact_pairs
  if (n == "contact_mpm_id") return plain(c_pid, sizeof(uint32_t) * n_contacts);
  if (n == "contact_rigid_id") return plain(c_rigid_id, sizeof(uint32_t) * n_contacts);
  if (n == "contact_dist") return plain(c_dist, sizeof(T) * n_contacts);
  if (n == "contact_normal") return plain(c_normal, sizeof(T) * 3 * n_contacts);
  if (n == "contact_pos") return plain(c_pos, sizeof(T) * 3 * n_contacts);
  if (n == "contact_rigid_v") return plain(c_rigid_v, sizeof(T) * 3 * n_contacts);
  if (n == "contact_rigid_p") return plain(c_rigid_p, sizeof(T) * 3 * n_contacts);
  throw std::out_of_range("gmpm_read: unknown buffer name '" + n + "'");
}

}  // namespace gmpm
