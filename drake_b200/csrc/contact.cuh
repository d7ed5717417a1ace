// contact.cuh -- MPM-rigid contact solve (replaces GpuMpmSolver<T>::CopyContactPairs / UpdateContact,
// cuda_mpm_solver.cu:194-621, and kernels K11-K17, K19 of cuda_mpm_kernels.cuh:604-630, 926-1658).
//
// The reference drives its Jacobi-Newton loop from the host: per Newton iteration ~10 small launches,
// 3-5 cudaMemsets and 4-8 blocking scalar read-backs; per line-search evaluation 2 launches and 3 more.
// Here the WHOLE solve is one cooperative persistent kernel (contact_newton_kernel): grid-wide
// barriers separate the phases, the convergence test and the line-search state machine
// (DoNewtonWithBisectionFallback restated, cuda_mpm_solver.cu:352-470, or backtracking, :472-527)
// run redundantly in every thread on device-resident scalars, and the host reads one stats record
// at the end.  Grid-side work is restricted to the blocks a contact stencil reaches ("contact
// blocks"): every other touched cell has Dir = 0 and v = v*, so its terms in the reference's sums are
// exact zeros and its Hess/Grad/Dir stay zero.  Hess/Grad/Dir live in a COMPACT per-contact-block
// store [slot][component][64] instead of the reference's three dense grids (120 MiB at fp32).
#pragma once
#include <cooperative_groups.h>

#include "state.cuh"

namespace gmpm {
namespace cg = cooperative_groups;

constexpr int kContactThreads = 256;
constexpr int kFrameComps = 14;  // mass, phi0, R_WC[9], v0_local[3]  (planar: frame[c * cap + i])

// Device-resident scalars of one solve.
template <typename T>
struct ContactScalars {
  T norm_acc;            // sum |Dir|^2 of the current Newton iteration (norm_dir_d, :252)
  T acc[3][4];           // rotating line-search accumulators {E0, E1, dE1, d2E1} (:253-259)
  uint32_t dofs;         // total_grid_DoFs_d (:260)
  uint32_t n_cblocks;    // number of contact blocks of this solve
  // results (gmpm_contact_stats)
  int iterations, grid_dofs, line_search_evals, pad;
  double residual, energy;
};

template <typename T>
struct ContactArgs {
  uint32_t n, cap;                         // contacts, planar stride of `frame`
  const T* pos;                            // [n][3] contact positions (host supplied)
  const T* rigid_v;                        // [n][3]
  const T* rigid_p;                        // [n][3]
  const uint32_t* rigid_id;                // [n]
  T* frame;                                // [kFrameComps][cap]
  T* vel;                                  // contact_vel  [n][3]
  T* vel0;                                 // contact_vel0 [n][3]
  const T* g_mass;                         // dense grid
  T* g_vel;                                // dense grid velocity (grid_momentum after UpdateGrid)
  const T* g_vstar;
  const uint32_t* cb_list;                 // contact blocks (grid block ids)
  const uint32_t* cb_slot;                 // grid block id -> slot in the compact store
  T *Hess, *Grad, *Dir;                    // compact [slot][9|3|3][64]
  T *F_tau, *F_f;                          // per-body impulses [n_bodies][3]
  uint32_t n_bodies;
  ContactScalars<T>* sc;
  T dt, mu, stiffness, damping, epsv;
  int exact, max_it;
  Consts<T> k;
};

// Quadratic B-spline stencil of a point (cuda_mpm_kernels.cuh:1067-1080 and the same lines in every
// transfer kernel).
template <typename T>
__device__ __forceinline__ void stencil(const T* x, const Consts<T>& k, uint32_t* base, T w[3][3]) {
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    base[d] = base_node(x[d], k.dx_inv, k.domain_size);
    const T fx = x[d] * k.dx_inv - T(base[d]);
    w[0][d] = T(0.5) * (T(1.5) - fx) * (T(1.5) - fx);
    w[1][d] = T(0.75) - (fx - T(1)) * (fx - T(1));
    w[2][d] = T(0.5) * (fx - T(0.5)) * (fx - T(0.5));
  }
}

// Lagged Hunt-Crossley local terms (compute_contact_grad_and_hess, cuda_mpm_kernels.cuh:956-1040):
// C_Grad = gamma (impulse), C_Hess = -G, in the contact frame (z = normal).
template <typename T>
__device__ __forceinline__ void contact_grad_hess(T epsv, T phi0, T dt, T stiffness, T damping, T mu, const T* v0,
                                                  const T* vn, T* H, T* G) {
  const T v_hat = tmin(phi0 / dt, T(1) / damping);
#pragma unroll
  for (int i = 0; i < 9; ++i) H[i] = T(0);
  G[0] = G[1] = G[2] = T(0);
  if (v0[2] > v_hat) return;
  const T yn = stiffness * dt * (phi0 - dt * vn[2]) * (T(1) - damping * vn[2]);
  const T d2 = stiffness * dt * (-dt - damping * phi0 + T(2) * damping * dt * vn[2]);
  const T yn0 = tmax(stiffness * dt * phi0 * (T(1) - damping * v0[2]), T(0));
  const T tsc = tsqrt(vn[0] * vn[0] + vn[1] * vn[1] + epsv * epsv);
  const T th0 = vn[0] / tsc, th1 = vn[1] / tsc;
  const T coeff = -mu * yn0 / tsc;
  H[0] = coeff * (T(1) - th0 * th0);
  H[1] = coeff * (-(th0 * th1));
  H[3] = coeff * (-(th1 * th0));
  H[4] = coeff * (T(1) - th1 * th1);
  H[8] = d2;
  G[0] = -mu * yn0 * th0;
  G[1] = -mu * yn0 * th1;
  G[2] = yn;
}

// Contact cost l(v) (the lambda at cuda_mpm_kernels.cuh:1423-1435).
template <typename T>
__device__ __forceinline__ T contact_cost(T epsv, T phi0, T dt, T stiffness, T damping, T mu, const T* v0, const T* v) {
  const T v_hat = tmin(phi0 / dt, T(1) / damping);
  const T yn0 = tmax(stiffness * dt * phi0 * (T(1) - damping * v0[2]), T(0));
  const T lt = mu * yn0 * (tsqrt(v[0] * v[0] + v[1] * v[1] + epsv * epsv) - epsv);
  const T vn = tmin(v_hat, v[2]);
  const T a = stiffness * damping * dt * dt;
  const T b = -(stiffness * dt * (dt + damping * phi0));
  const T c = stiffness * dt * phi0;
  const T ln = -(T(1. / 3.) * a * vn * vn * vn + T(1. / 2.) * b * vn * vn + c * vn);
  return lt + ln;
}

// initialize_contact_velocities (:926-938): contact_vel[i] = velocity of the particle in contact.
template <typename T>
__global__ void contact_init_vel_kernel(uint32_t n, uint32_t N, const uint32_t* __restrict__ mpm_id,
                                        const int* __restrict__ pid_of_vis, const int* __restrict__ slot_of_pid,
                                        const T* __restrict__ vel, T* __restrict__ c_vel) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = slot_of_pid[pid_of_vis[mpm_id[i]]];
#pragma unroll
  for (int d = 0; d < 3; ++d) c_vel[i * 3 + d] = vel[(size_t)d * N + s];
}

// Gated velocity gather of one contact point (grid_to_particle_kernel<.., CONTACT_TRANSFER=true>, :798-924).
template <typename T>
__device__ __forceinline__ void contact_gather(const T* x, const Consts<T>& k, const T* g_mass, const T* g_vel, T* out) {
  uint32_t base[3];
  T w[3][3];
  stencil(x, k, base, w);
  T v[3] = {T(0), T(0), T(0)};
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int kk = 0; kk < 3; ++kk) {
        const uint32_t key = cell_key(base[0] + i, base[1] + j, base[2] + kk);
        const T wt = w[i][0] * w[j][1] * w[kk][2];
        if (g_mass[key] > T(1e-7)) {
          const size_t vb = (size_t)(key >> 6) * 192 + (key & 63u);
          v[0] += wt * g_vel[vb];
          v[1] += wt * g_vel[vb + 64];
          v[2] += wt * g_vel[vb + 128];
        }
      }
  out[0] = v[0];
  out[1] = v[1];
  out[2] = v[2];
}

// Per-solve setup, one thread per contact: sort key of the contact point (:228-232), the contact
// frame and lagged quantities shared by K13 and K15 (:1107-1142, :1386-1423), contact_vel0 (:267-272),
// and the list of contact blocks with their compact slots.
template <typename T>
__global__ void contact_setup_kernel(uint32_t n, uint32_t cap, uint32_t N, const uint32_t* __restrict__ mpm_id,
                                     const int* __restrict__ pid_of_vis, const int* __restrict__ slot_of_pid,
                                     const T* __restrict__ vel, const T* __restrict__ vol,
                                     const T* __restrict__ c_pos, const T* __restrict__ c_dist,
                                     const T* __restrict__ c_normal, const T* __restrict__ c_rigid_v,
                                     const T* __restrict__ g_mass, const T* __restrict__ g_vel,
                                     uint32_t* __restrict__ c_keys, T* __restrict__ frame, T* __restrict__ c_vel0,
                                     uint32_t* cb_epoch, uint32_t* cb_slot, uint32_t* cb_list, uint32_t epoch,
                                     uint32_t cb_cap, ContactScalars<T>* sc, Consts<T> k) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const T x[3] = {c_pos[i * 3], c_pos[i * 3 + 1], c_pos[i * 3 + 2]};
  uint32_t b[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) b[d] = base_node(x[d], k.dx_inv, k.domain_size);
  c_keys[i] = cell_key(b[0], b[1], b[2]);
  const int s = slot_of_pid[pid_of_vis[mpm_id[i]]];
  const T mass = vol[s] * k.density;
  const T nhat[3] = {-c_normal[i * 3], -c_normal[i * 3 + 1], -c_normal[i * 3 + 2]};
  T R[9];
  frame_from_unit_vector(nhat, 2, R);
  const T v0r[3] = {vel[s] - c_rigid_v[i * 3], vel[(size_t)N + s] - c_rigid_v[i * 3 + 1],
                    vel[(size_t)2 * N + s] - c_rigid_v[i * 3 + 2]};
  T v0l[3];
  mul33v(R, v0r, v0l);
  frame[i] = mass;
  frame[(size_t)cap + i] = -c_dist[i];
#pragma unroll
  for (int e = 0; e < 9; ++e) frame[(size_t)(2 + e) * cap + i] = R[e];
#pragma unroll
  for (int d = 0; d < 3; ++d) frame[(size_t)(11 + d) * cap + i] = v0l[d];
  T v0[3];
  contact_gather(x, k, g_mass, g_vel, v0);
  c_vel0[i * 3] = v0[0];
  c_vel0[i * 3 + 1] = v0[1];
  c_vel0[i * 3 + 2] = v0[2];
  // contact blocks: the stencil base..base+2 reaches at most 2 blocks per axis
  const uint32_t bpa = (uint32_t)k.domain_size >> 2;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const uint32_t ox = (q >> 2) & 1u, oy = (q >> 1) & 1u, oz = q & 1u;
    if ((ox && (b[0] & 3u) < 2u) || (oy && (b[1] & 3u) < 2u) || (oz && (b[2] & 3u) < 2u)) continue;
    const uint32_t gx = (b[0] >> 2) + ox, gy = (b[1] >> 2) + oy, gz = (b[2] >> 2) + oz;
    if (gx >= bpa || gy >= bpa || gz >= bpa) continue;
    const uint32_t g = morton3(gx, gy, gz);
    if (atomicExch(&cb_epoch[g], epoch) != epoch) {
      const uint32_t slot = atomicAdd(&sc->n_cblocks, 1u);
      if (slot < cb_cap) {
        cb_list[slot] = g;
        cb_slot[g] = slot;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// The Newton solve: one cooperative persistent kernel.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(kFullMask, v, o);
  return v;
}
template <typename T>
__device__ __forceinline__ int tsign(T v) { return v > T(0) ? 1 : (v < T(0) ? -1 : 0); }

template <typename T>
struct ContactFrame {
  T mass, phi0, R[9], v0[3];
};
template <typename T>
__device__ __forceinline__ void load_frame(const ContactArgs<T>& a, uint32_t i, ContactFrame<T>& f) {
  f.mass = a.frame[i];
  f.phi0 = a.frame[(size_t)a.cap + i];
#pragma unroll
  for (int e = 0; e < 9; ++e) f.R[e] = a.frame[(size_t)(2 + e) * a.cap + i];
#pragma unroll
  for (int d = 0; d < 3; ++d) f.v0[d] = a.frame[(size_t)(11 + d) * a.cap + i];
}

// K12 clean_grid_contact_kernel (:604-630) on the compact store.
template <typename T>
__device__ __forceinline__ void phase_clean(const ContactArgs<T>& a, uint32_t n_cb, uint32_t tid, uint32_t nth) {
  const size_t nh = (size_t)n_cb * 9 * 64, ng = (size_t)n_cb * 3 * 64;
  for (size_t i = tid; i < nh; i += nth) a.Hess[i] = T(0);
  for (size_t i = tid; i < ng; i += nth) {
    a.Grad[i] = T(0);
    a.Dir[i] = T(0);
  }
  if (tid == 0) {
    a.sc->norm_acc = T(0);
    a.sc->dofs = 0u;
  }
}

// K13 contact_particle_to_grid_kernel<JACOBI=true> (:1042-1215): per-contact Jacobian row (R_CW w_ip,
// implicit), local impulse and Hessian block, rotated to world and lumped onto the 27 nodes.
//
// Warp merge before the scatter (the reference's scheme, :1077-1083, :1169, :1196, same as its P2G :448-523):
// the lanes of a warp hold CONSECUTIVE contacts, and consecutive contacts are particles that are neighbours in the
// visible order (cell order after a sort, mesh rows before), so runs of lanes share a base cell and with it all 27
// destination nodes.  The 12 Hess/Grad scalars of a node are summed over each run with a segmented shuffle
// reduction and only the run's first lane issues the atomics.  Unlike upstream the contacts are not re-sorted
// for this (the merge takes the runs the pair order already has); a warp without any run skips the shuffles.
template <typename T>
__device__ __forceinline__ void phase_contact_p2g(const ContactArgs<T>& a, uint32_t tid, uint32_t nth) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t rounds = (a.n + nth - 1) / nth;  // uniform trip count: the shuffles need all lanes
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t i = r * nth + tid;
    const bool valid = i < a.n;
    uint32_t base[3] = {0u, 0u, 0u};
    T w[3][3];
    T WG[3], WH[9];
    T mass = T(0);
#pragma unroll
    for (int q = 0; q < 3; ++q) w[q][0] = w[q][1] = w[q][2] = T(0);
#pragma unroll
    for (int e = 0; e < 9; ++e) WH[e] = T(0);
    WG[0] = WG[1] = WG[2] = T(0);
    if (valid) {
      const T x[3] = {a.pos[i * 3], a.pos[i * 3 + 1], a.pos[i * 3 + 2]};
      stencil(x, a.k, base, w);
      ContactFrame<T> f;
      load_frame(a, i, f);
      mass = f.mass;
      const T vr[3] = {a.vel[i * 3] - a.rigid_v[i * 3], a.vel[i * 3 + 1] - a.rigid_v[i * 3 + 1],
                       a.vel[i * 3 + 2] - a.rigid_v[i * 3 + 2]};
      T vn[3], CH[9], CG[3];
      mul33v(f.R, vr, vn);
      contact_grad_hess(a.epsv, f.phi0, a.dt, a.stiffness, a.damping, a.mu, f.v0, vn, CH, CG);
      T tmp[9], Rt[9];
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int c = 0; c < 3; ++c) Rt[rr * 3 + c] = f.R[c * 3 + rr];
      mul33v(Rt, CG, WG);
      mul33(Rt, CH, tmp);
      mul33(tmp, f.R, WH);
    }
    // runs of equal base cell among the lanes (a lane without a contact is a run of its own)
    const uint32_t mykey = valid ? cell_key(base[0], base[1], base[2]) : (0xFFFFFFFFu - lane);
    const uint32_t prevkey = __shfl_up_sync(kFullMask, mykey, 1);
    const bool head = lane == 0u || prevkey != mykey;
    const uint32_t heads = __ballot_sync(kFullMask, head);
    const bool merge = heads != kFullMask;                                   // warp-uniform
    const uint32_t above = lane == 31u ? 0u : (heads & ~((2u << lane) - 1u));  // run heads after this lane
    const uint32_t span = (above ? (uint32_t)__ffs((int)above) - 1u : 32u) - lane;  // lanes from here to the run's end
#pragma unroll
    for (int ii = 0; ii < 3; ++ii)
#pragma unroll
      for (int jj = 0; jj < 3; ++jj)
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) {
          const T wt = w[ii][0] * w[jj][1] * w[kk][2];
          T val[12];
#pragma unroll
          for (int e = 0; e < 9; ++e) val[e] = mass * wt * wt * WH[e];
#pragma unroll
          for (int d = 0; d < 3; ++d) val[9 + d] = mass * wt * WG[d];
          if (merge) {
#pragma unroll
            for (int e = 0; e < 12; ++e)
#pragma unroll
              for (uint32_t o = 1; o < 32u; o <<= 1) {
                const T t = __shfl_down_sync(kFullMask, val[e], o);
                if (o < span) val[e] += t;
              }
          }
          if (valid && head) {
            const uint32_t key = cell_key(base[0] + ii, base[1] + jj, base[2] + kk);
            const size_t slot = a.cb_slot[key >> 6];
            const uint32_t l = key & 63u;
            T* H = a.Hess + slot * (9 * 64) + l;
            T* G = a.Grad + slot * (3 * 64) + l;
#pragma unroll
            for (int e = 0; e < 9; ++e) atomicAdd(H + e * 64, val[e]);
#pragma unroll
            for (int d = 0; d < 3; ++d) atomicAdd(G + d * 64, val[9 + d]);
          }
        }
  }
}

// K14 update_grid_contact_coordinate_descent_kernel<JACOBI=true> (:1217-1274).
template <typename T>
__device__ __forceinline__ void phase_newton_dir(const ContactArgs<T>& a, uint32_t n_cb, uint32_t tid, uint32_t nth,
                                                 T relax) {
  T nsum = T(0);
  uint32_t dofs = 0;
  const uint32_t total = n_cb * 64u;
  const uint32_t rounds = (total + nth - 1) / nth;  // uniform trip count: the warp reduction needs all lanes
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t c = r * nth + tid;
    if (c < total) {
      const uint32_t slot = c >> 6, l = c & 63u;
      const uint32_t g = a.cb_list[slot];
      const T m = a.g_mass[(size_t)g * 64 + l];
      T* Hp = a.Hess + (size_t)slot * (9 * 64) + l;
      T* Gp = a.Grad + (size_t)slot * (3 * 64) + l;
      T H[9], G[3];
      T hn = T(0), gn = T(0);
#pragma unroll
      for (int e = 0; e < 9; ++e) {
        H[e] = Hp[e * 64];
        hn += H[e] * H[e];
      }
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        G[d] = Gp[d * 64];
        gn += G[d] * G[d];
      }
      if (m > T(0) && ((double)tsqrt(hn) > 1e-7 || (double)tsqrt(gn) > 1e-7)) {
        const size_t vb = (size_t)g * 192 + l;
        H[0] -= m;
        H[4] -= m;
        H[8] -= m;
#pragma unroll
        for (int d = 0; d < 3; ++d) G[d] -= m * (a.g_vel[vb + d * 64] - a.g_vstar[vb + d * 64]);
        T Hi[9], D[3];
        inverse3(H, Hi);
        mul33v(Hi, G, D);
        nsum += D[0] * D[0] + D[1] * D[1] + D[2] * D[2];
        dofs += 1u;
#pragma unroll
        for (int e = 0; e < 9; ++e) Hp[e * 64] = H[e];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          Gp[d * 64] = G[d];
          a.Dir[(size_t)slot * (3 * 64) + d * 64 + l] = D[d] * relax;
        }
      }
    }
  }
  nsum = warp_sum(nsum);
  dofs = warp_sum(dofs);
  if ((threadIdx.x & 31) == 0 && dofs) {
    atomicAdd(&a.sc->norm_acc, nsum);
    atomicAdd(&a.sc->dofs, dofs);
  }
}

// K15 grid_to_particle_vdb_line_search_kernel<JACOBI=true> (:1276-1473) + K16
// update_global_energy_grid_kernel (:1536-1589): E(alpha) and optionally E(0), dE/dalpha, d2E/dalpha2.
template <typename T>
__device__ __forceinline__ void phase_energy(const ContactArgs<T>& a, uint32_t n_cb, uint32_t tid, uint32_t nth,
                                             T alpha, bool eval_E0, bool solve_df, T* acc) {
  T e0 = T(0), e1 = T(0), de = T(0), d2e = T(0);
  for (uint32_t i = tid; i < a.n; i += nth) {
    const T x[3] = {a.pos[i * 3], a.pos[i * 3 + 1], a.pos[i * 3 + 2]};
    uint32_t base[3];
    T w[3][3];
    stencil(x, a.k, base, w);
    T ov[3] = {T(0), T(0), T(0)}, nv[3] = {T(0), T(0), T(0)}, gd[3] = {T(0), T(0), T(0)};
#pragma unroll
    for (int ii = 0; ii < 3; ++ii)
#pragma unroll
      for (int jj = 0; jj < 3; ++jj)
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) {
          const T wt = w[ii][0] * w[jj][1] * w[kk][2];
          const uint32_t key = cell_key(base[0] + ii, base[1] + jj, base[2] + kk);
          const uint32_t l = key & 63u;
          const size_t vb = (size_t)(key >> 6) * 192 + l;
          const size_t db = (size_t)a.cb_slot[key >> 6] * 192 + l;
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            const T gv = a.g_vel[vb + d * 64], gD = a.Dir[db + d * 64];
            ov[d] += wt * gv;
            nv[d] += wt * (gv - alpha * gD);
            gd[d] += wt * gD;
          }
        }
    ContactFrame<T> f;
    load_frame(a, i, f);
    T orl[3], nrl[3], ol[3], nl[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      orl[d] = ov[d] - a.rigid_v[i * 3 + d];
      nrl[d] = nv[d] - a.rigid_v[i * 3 + d];
    }
    mul33v(f.R, orl, ol);
    mul33v(f.R, nrl, nl);
    e1 += f.mass * contact_cost(a.epsv, f.phi0, a.dt, a.stiffness, a.damping, a.mu, f.v0, nl);
    if (solve_df) {
      T CH[9], CG[3], RD[3], t[3];
      contact_grad_hess(a.epsv, f.phi0, a.dt, a.stiffness, a.damping, a.mu, f.v0, nl, CH, CG);
      mul33v(f.R, gd, RD);
      de += f.mass * (CG[0] * RD[0] + CG[1] * RD[1] + CG[2] * RD[2]);
      mul33Tv(CH, RD, t);  // RD^T C_Hess
      d2e += f.mass * (t[0] * RD[0] + t[1] * RD[1] + t[2] * RD[2]);
    }
    if (eval_E0) e0 += f.mass * contact_cost(a.epsv, f.phi0, a.dt, a.stiffness, a.damping, a.mu, f.v0, ol);
  }
  const uint32_t total = n_cb * 64u;
  for (uint32_t c = tid; c < total; c += nth) {
    const uint32_t slot = c >> 6, l = c & 63u;
    const uint32_t g = a.cb_list[slot];
    const T m = a.g_mass[(size_t)g * 64 + l];
    if (m > T(0)) {
      const size_t vb = (size_t)g * 192 + l, db = (size_t)slot * 192 + l;
      T o2 = T(0), n2 = T(0), nd = T(0), dd = T(0);
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const T gv = a.g_vel[vb + d * 64], vs = a.g_vstar[vb + d * 64], D = a.Dir[db + d * 64];
        const T o = gv - vs, nn = gv - alpha * D - vs;
        o2 += o * o;
        n2 += nn * nn;
        nd += nn * D;
        dd += D * D;
      }
      if (eval_E0) e0 += T(0.5) * m * o2;
      e1 += T(0.5) * m * n2;
      if (solve_df) {
        de += -m * nd;
        d2e += m * dd;
      }
    }
  }
  // CTA reduction, then one atomic per CTA and quantity
  __shared__ T red[4][kContactThreads / 32];
  e0 = warp_sum(e0);
  e1 = warp_sum(e1);
  de = warp_sum(de);
  d2e = warp_sum(d2e);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    red[0][warp] = e0;
    red[1][warp] = e1;
    red[2][warp] = de;
    red[3][warp] = d2e;
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      T v = lane < kContactThreads / 32 ? red[q][lane] : T(0);
      v = warp_sum(v);
      if (lane == 0 && (q == 1 || (q == 0 && eval_E0) || (q >= 2 && solve_df))) atomicAdd(&acc[q], v);
    }
  }
  __syncthreads();
}

// K17 apply_global_line_search_grid_kernel (:1591-1614).
template <typename T>
__device__ __forceinline__ void phase_apply(const ContactArgs<T>& a, uint32_t n_cb, uint32_t tid, uint32_t nth,
                                            T alpha) {
  const uint32_t total = n_cb * 64u;
  for (uint32_t c = tid; c < total; c += nth) {
    const uint32_t slot = c >> 6, l = c & 63u;
    const uint32_t g = a.cb_list[slot];
    if (a.g_mass[(size_t)g * 64 + l] > T(0)) {
      const size_t vb = (size_t)g * 192 + l, db = (size_t)slot * 192 + l;
#pragma unroll
      for (int d = 0; d < 3; ++d) a.g_vel[vb + d * 64] -= alpha * a.Dir[db + d * 64];
    }
  }
}

// K19 apply_contact_impulse_to_rigid_bodies (:1616-1658).  Per-CTA shared accumulators for the first
// 32 bodies (the reference issues 6 global atomics per contact onto a handful of addresses).
template <typename T>
__device__ __forceinline__ void phase_impulses(const ContactArgs<T>& a, uint32_t tid, uint32_t nth) {
  constexpr int kSharedBodies = 32;
  __shared__ T sacc[kSharedBodies][6];
  for (int i = threadIdx.x; i < kSharedBodies * 6; i += blockDim.x) (&sacc[0][0])[i] = T(0);
  __syncthreads();
  for (uint32_t i = tid; i < a.n; i += nth) {
    const uint32_t body = a.rigid_id[i];
    if (body >= a.n_bodies) continue;
    const T m = a.frame[i];
    T l[3], p[3], h[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      l[d] = m * -(a.vel[i * 3 + d] - a.vel0[i * 3 + d]);
      p[d] = a.pos[i * 3 + d] - a.rigid_p[i * 3 + d];
    }
    h[0] = p[1] * l[2] - l[1] * p[2];
    h[1] = p[2] * l[0] - l[2] * p[0];
    h[2] = p[0] * l[1] - l[0] * p[1];
    if (body < (uint32_t)kSharedBodies) {
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        atomicAdd(&sacc[body][d], h[d]);
        atomicAdd(&sacc[body][3 + d], l[d]);
      }
    } else {
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        atomicAdd(&a.F_tau[(size_t)body * 3 + d], h[d]);
        atomicAdd(&a.F_f[(size_t)body * 3 + d], l[d]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSharedBodies * 6; i += blockDim.x) {
    const uint32_t body = i / 6, c = i % 6;
    const T v = sacc[body][c];
    if (body < a.n_bodies && v != T(0)) atomicAdd(c < 3 ? &a.F_tau[body * 3 + c] : &a.F_f[body * 3 + (c - 3)], v);
  }
}

template <typename T>
__device__ __forceinline__ T ld_scalar(const T* p) { return __ldcg(p); }

// The solve (cuda_mpm_solver.cu:215-621, use_jacobi = true, global line search).  Every thread runs
// the same scalar control flow on the same device scalars, so all grid barriers are reached uniformly.
template <typename T>
__global__ void __launch_bounds__(kContactThreads)
contact_newton_kernel(ContactArgs<T> a) {
  cg::grid_group grid = cg::this_grid();
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t nth = gridDim.x * blockDim.x;
  ContactScalars<T>* sc = a.sc;
  const uint32_t n_cb = __ldcg(&sc->n_cblocks);
  const T kTol = T(1e-4), relax = T(0.3);                 // :236, :239
  const T f_tol = T(1e-8), x_tol = f_tol * relax;         // :362-364
  const int kMaxLineSearch = 200;                         // safety bound (the reference loops unbounded)
  int count = 0, grid_dofs = 0, ls_total = 0;
  T norm_dir = T(1e10), E1g = T(0);

  phase_clean(a, n_cb, tid, nth);
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) sc->acc[0][q] = T(0);
  }
  grid.sync();
  while (norm_dir > kTol && count < a.max_it) {
    phase_contact_p2g(a, tid, nth);
    grid.sync();
    phase_newton_dir(a, n_cb, tid, nth, relax);
    grid.sync();
    const T norm_acc = ld_scalar(&sc->norm_acc);
    const uint32_t total_dofs = __ldcg(&sc->dofs);

    // ---- line search ----------------------------------------------------------------------------
    int ev = 0;  // evaluations issued so far in this Newton iteration (accumulator slot = ev % 3)
    auto eval = [&](T alpha, bool e0, bool df, T* out) {
      const int slot = ev % 3, nxt = (ev + 1) % 3;
      if (tid == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sc->acc[nxt][q] = T(0);
      }
      phase_energy(a, n_cb, tid, nth, alpha, e0, df, &sc->acc[slot][0]);
      grid.sync();
#pragma unroll
      for (int q = 0; q < 4; ++q) out[q] = ld_scalar(&sc->acc[slot][q]);
      ++ev;
    };
    T alpha = T(1);
    int ls_cnt = 0;
    if (a.exact) {
      T fl[4] = {T(0), T(0), T(0), T(0)}, fu[4], fr[4];
      T x_lower = T(0), x_upper = T(1), root;
      eval(T(0), false, true, fl);
      eval(T(1), false, true, fu);
      if (fl[2] < T(0) && fu[2] < T(0)) {  // :393-396
        x_lower = T(1);
#pragma unroll
        for (int q = 0; q < 4; ++q) fl[q] = fu[q];
      }
      root = x_upper;
      T minus_dx = x_lower - x_upper, minus_dx_prev = minus_dx;
      bool ok = false;
      while (!ok && ls_cnt < kMaxLineSearch) {
        eval(root, false, true, fr);
        if (tsign(fr[2]) != tsign(fu[2])) {
          x_lower = root;
        } else {
          x_upper = root;
#pragma unroll
          for (int q = 0; q < 4; ++q) fu[q] = fr[q];
        }
        if (tabs(fr[2]) < f_tol) ok = true;
        const bool newton_is_slow = T(2) * tabs(fr[2]) > tabs(minus_dx_prev * fr[3]);
        minus_dx_prev = minus_dx;
        if (newton_is_slow) {
          const T dxn = T(.5) * (x_lower - x_upper);
          root = x_lower - dxn;
          minus_dx = dxn;
        } else {
          const T dxn = fr[2] / fr[3];
          T x = root;
          x -= dxn;
          root = x;
          minus_dx = dxn;
          if (!(x_lower <= root && root <= x_upper)) {
            const T dxb = T(.5) * (x_lower - x_upper);
            root = x_lower - dxb;
            minus_dx = dxb;
          }
        }
        if (tabs(minus_dx) < x_tol) ok = true;
        if (ok) {
          E1g = fr[1];
          alpha = root;
        }
        ++ls_cnt;
      }
      if (!ok) {  // bound hit: take the last evaluated root
        E1g = fr[1];
        alpha = root;
      }
    } else {
      T E0g = T(0);
      bool ok = false;
      while (!ok && ls_cnt < kMaxLineSearch) {
        T f[4];
        eval(alpha, ls_cnt == 0, false, f);
        if (ls_cnt == 0) E0g = f[0];
        E1g = f[1];
        if (E1g <= E0g) {
          ok = true;
        } else {
          alpha /= T(2);
          if (alpha < T(1e-8)) ok = true;
        }
        ++ls_cnt;
      }
    }
    ls_total += ls_cnt;
    phase_apply(a, n_cb, tid, nth, alpha);
    grid.sync();
    // contact_vel <- gated gather (:557-562); merged with the clean of the next iteration
    for (uint32_t i = tid; i < a.n; i += nth) {
      const T x[3] = {a.pos[i * 3], a.pos[i * 3 + 1], a.pos[i * 3 + 2]};
      T v[3];
      contact_gather(x, a.k, a.g_mass, a.g_vel, v);
      a.vel[i * 3] = v[0];
      a.vel[i * 3 + 1] = v[1];
      a.vel[i * 3 + 2] = v[2];
    }
    grid_dofs = (int)total_dofs;
    norm_dir = tsqrt(norm_acc) / T(grid_dofs);  // :569
    count += 1;
    if (norm_dir > kTol && count < a.max_it) {
      phase_clean(a, n_cb, tid, nth);
      if (tid == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) sc->acc[0][q] = T(0);
      }
    }
    grid.sync();
  }
  phase_impulses(a, tid, nth);
  if (tid == 0) {
    sc->iterations = count;
    sc->grid_dofs = grid_dofs;
    sc->line_search_evals = ls_total;
    sc->residual = (double)norm_dir;
    sc->energy = (double)E1g;
  }
}

// Getter support: scatter one compact contact-grid field to the reference's dense layout [cell][comps].
template <typename T>
__global__ void contact_grid_to_reference_kernel(uint32_t n_cb, int comps, const uint32_t* __restrict__ cb_list,
                                                 const T* __restrict__ src, T* __restrict__ out) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cb * 64u) return;
  const uint32_t slot = c >> 6, l = c & 63u;
  const size_t cell = (size_t)cb_list[slot] * 64 + l;
  for (int e = 0; e < comps; ++e) out[cell * comps + e] = src[((size_t)slot * comps + e) * 64 + l];
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
template <typename T>
void State<T>::dfree(void* p) {
  if (!p) return;
  auto it = std::find(allocs.begin(), allocs.end(), p);
  if (it != allocs.end()) allocs.erase(it);
  cudaFree(p);
}

// GpuMpmState<T>::ReallocateContacts (cuda_mpm_model.cu:268-317): grow-only.
template <typename T>
void State<T>::reallocate_contacts(size_t n) {
  if (n <= contact_cap) return;
  GMPM_CUDA(cudaStreamSynchronize(stream));
  const size_t cap = std::max<size_t>(n + n / 2, 1024);
  for (void* p : {(void*)c_pid, (void*)c_rigid_id, (void*)c_keys, (void*)c_pos, (void*)c_vel, (void*)c_vel0,
                  (void*)c_dist, (void*)c_normal, (void*)c_rigid_v, (void*)c_rigid_p, (void*)c_frame,
                  (void*)cb_list, (void*)g_Hess, (void*)g_Grad, (void*)g_Dir})
    dfree(p);
  c_pid = dalloc<uint32_t>(cap);
  c_rigid_id = dalloc<uint32_t>(cap);
  c_keys = dalloc<uint32_t>(cap);
  c_pos = dalloc<T>(3 * cap);
  c_vel = dalloc<T>(3 * cap);
  c_vel0 = dalloc<T>(3 * cap);
  c_dist = dalloc<T>(cap);
  c_normal = dalloc<T>(3 * cap);
  c_rigid_v = dalloc<T>(3 * cap);
  c_rigid_p = dalloc<T>(3 * cap);
  c_frame = dalloc<T>((size_t)kFrameComps * cap);
  cb_cap = std::min<size_t>(k.n_blocks, 8 * cap);
  cb_list = dalloc<uint32_t>(cb_cap);
  g_Hess = dalloc<T>(cb_cap * 9 * 64);
  g_Grad = dalloc<T>(cb_cap * 3 * 64);
  g_Dir = dalloc<T>(cb_cap * 3 * 64);
  if (!cb_epoch) {
    cb_epoch = dalloc<uint32_t>(k.n_blocks);
    cb_slot = dalloc<uint32_t>(k.n_blocks);
    csc = dalloc<ContactScalars<T>>(1);
    GMPM_CUDA(cudaMallocHost(&h_csc, sizeof(ContactScalars<T>)));
  }
  contact_cap = cap;
  // pinned staging for the 7 host arrays of one CopyContactPairs call
  const size_t stage = cap * (2 * sizeof(uint32_t) + 13 * sizeof(T));
  if (stage > h_stage_cap) {
    if (h_stage) cudaFreeHost(h_stage);
    h_stage = nullptr;
    GMPM_CUDA(cudaMallocHost(&h_stage, stage));
    h_stage_cap = stage;
  }
}

// GpuMpmSolver<T>::CopyContactPairs (cuda_mpm_solver.cu:194-212).  The host arrays are borrowed
// (pageable): they are packed into one pinned staging buffer and uploaded asynchronously.
template <typename T>
void State<T>::copy_contact_pairs(size_t n, const uint32_t* mpm_id, const uint32_t* rigid_id, const T* dist,
                                  const T* normal, const T* cpos, const T* rigid_v, const T* rigid_p) {
  require_finalized();
  if (n >= (size_t)1 << 31) throw std::invalid_argument("too many contact pairs");
  cb_valid = false;
  if (n == 0) {
    n_contacts = 0;
    return;
  }
  if (!mpm_id || !rigid_id || !dist || !normal || !cpos || !rigid_v || !rigid_p)
    throw std::invalid_argument("CopyContactPairs: null array");
  for (size_t i = 0; i < n; ++i)
    if (mpm_id[i] >= N) throw std::invalid_argument("CopyContactPairs: particle index out of range");
  reallocate_contacts(n);
  GMPM_CUDA(cudaStreamSynchronize(stream));  // the staging buffer of the previous call is free again
  n_contacts = n;
  char* h = static_cast<char*>(h_stage);
  struct Item { void* dst; const void* src; size_t bytes; };
  const Item items[7] = {{c_pid, mpm_id, n * sizeof(uint32_t)},     {c_rigid_id, rigid_id, n * sizeof(uint32_t)},
                         {c_pos, cpos, 3 * n * sizeof(T)},          {c_dist, dist, n * sizeof(T)},
                         {c_normal, normal, 3 * n * sizeof(T)},     {c_rigid_v, rigid_v, 3 * n * sizeof(T)},
                         {c_rigid_p, rigid_p, 3 * n * sizeof(T)}};
  size_t off = 0;
  note("contact_h2d");
  for (const Item& it : items) {
    std::memcpy(h + off, it.src, it.bytes);
    GMPM_CUDA(cudaMemcpyAsync(it.dst, h + off, it.bytes, cudaMemcpyHostToDevice, stream));
    off += it.bytes;
  }
  launch("contact_init_vel");
  contact_init_vel_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>((uint32_t)n, (uint32_t)N, c_pid,
                                                                    pid_of_vis[vis_cur], slot_of_pid, vel[cur], c_vel);
  GMPM_CUDA(cudaGetLastError());
}

template <typename T>
void State<T>::reallocate_external_bodies(size_t n) {  // cuda_mpm_model.cu:320-338
  bind_device();
  if (!stream) throw std::invalid_argument("state not created");
  if (n > body_cap) {
    GMPM_CUDA(cudaStreamSynchronize(stream));
    dfree(F_tau);
    dfree(F_f);
    F_tau = dalloc<T>(3 * n);
    F_f = dalloc<T>(3 * n);
    body_cap = n;
  }
  n_bodies = n;
  if (body_cap) {
    note("memset");
    GMPM_CUDA(cudaMemsetAsync(F_tau, 0, sizeof(T) * 3 * body_cap, stream));
    GMPM_CUDA(cudaMemsetAsync(F_f, 0, sizeof(T) * 3 * body_cap, stream));
  }
}

template <typename T>
void State<T>::external_body_force_to_host(T* tau, T* f) {  // cuda_mpm_model.cu:341-345
  bind_device();
  if (!n_bodies) return;
  GMPM_CUDA(cudaMemcpyAsync(tau, F_tau, sizeof(T) * 3 * n_bodies, cudaMemcpyDeviceToHost, stream));
  to_host(f, F_f, sizeof(T) * 3 * n_bodies);
}

// GpuMpmSolver<T>::UpdateContact (cuda_mpm_solver.cu:215-621).
template <typename T>
void State<T>::update_contact(T dt, T mu, T stiffness, T damping, bool exact, int max_it, gmpm_contact_stats* stats) {
  require_finalized();
  if (stats) std::memset(stats, 0, sizeof(*stats));
  if (!n_contacts) return;  // :217
  if (stage != GridStage::kVelocity)
    throw std::invalid_argument("UpdateContact needs the grid velocities: call UpdateGrid first");
  const uint32_t n = (uint32_t)n_contacts;
  ++contact_epoch;
  note("memset");
  GMPM_CUDA(cudaMemsetAsync(csc, 0, sizeof(ContactScalars<T>), stream));
  launch("contact_setup");
  contact_setup_kernel<T><<<blocks_for(n), kThreads, 0, stream>>>(
      n, (uint32_t)contact_cap, (uint32_t)N, c_pid, pid_of_vis[vis_cur], slot_of_pid, vel[cur], vol[cur], c_pos,
      c_dist, c_normal, c_rigid_v, g_mass, g_vel, c_keys, c_frame, c_vel0, cb_epoch, cb_slot, cb_list, contact_epoch,
      (uint32_t)cb_cap, csc, k);
  GMPM_CUDA(cudaGetLastError());
  cb_valid = true;

  ContactArgs<T> a{};
  a.n = n;
  a.cap = (uint32_t)contact_cap;
  a.pos = c_pos;
  a.rigid_v = c_rigid_v;
  a.rigid_p = c_rigid_p;
  a.rigid_id = c_rigid_id;
  a.frame = c_frame;
  a.vel = c_vel;
  a.vel0 = c_vel0;
  a.g_mass = g_mass;
  a.g_vel = g_vel;
  a.g_vstar = g_vstar;
  a.cb_list = cb_list;
  a.cb_slot = cb_slot;
  a.Hess = g_Hess;
  a.Grad = g_Grad;
  a.Dir = g_Dir;
  a.F_tau = F_tau;
  a.F_f = F_f;
  a.n_bodies = (uint32_t)n_bodies;
  a.sc = csc;
  a.dt = dt;
  a.mu = mu;
  a.stiffness = stiffness;
  a.damping = damping;
  a.epsv = k.epsv;
  a.exact = exact ? 1 : 0;
  a.max_it = max_it > 0 ? max_it : 2000;  // :234
  a.k = k;
  if (newton_ctas == 0) {
    int per_sm = 0;
    GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, contact_newton_kernel<T>, kContactThreads, 0));
    if (per_sm < 1) throw CudaError(cudaErrorLaunchOutOfResources, "contact_newton_kernel does not fit on an SM");
    newton_ctas = sm_count * std::min(per_sm, 2);
  }
  // enough CTAs for the work, never more than can be co-resident (cooperative launch)
  const size_t work = std::max<size_t>(n, std::min<size_t>(cb_cap, (size_t)8 * n) * 64);
  const int ctas = (int)std::max<size_t>(1, std::min<size_t>((size_t)newton_ctas, (work + kContactThreads - 1) / kContactThreads));
  void* kargs[] = {&a};
  launch("contact_newton");
  GMPM_CUDA(cudaLaunchCooperativeKernel((void*)contact_newton_kernel<T>, dim3(ctas), dim3(kContactThreads), kargs, 0,
                                        stream));
  if (stats) {
    GMPM_CUDA(cudaMemcpyAsync(h_csc, csc, sizeof(ContactScalars<T>), cudaMemcpyDeviceToHost, stream));
    GMPM_CUDA(cudaStreamSynchronize(stream));
    const ContactScalars<T>* r = static_cast<const ContactScalars<T>*>(h_csc);
    stats->iterations = r->iterations;
    stats->grid_dofs = r->grid_dofs;
    stats->line_search_evals = r->line_search_evals;
    stats->residual = r->residual;
    stats->energy = r->energy;
    if (prof.on) prof.flush(stream);
  }
}

// Dense reference-layout copy of a compact contact-grid field into the io buffer (getters).
template <typename T>
T* State<T>::contact_grid_dense(const T* src, int comps) {
  const size_t bytes = sizeof(T) * comps * k.n_cells;
  T* buf = static_cast<T*>(io(bytes));
  GMPM_CUDA(cudaMemsetAsync(buf, 0, bytes, stream));
  if (cb_valid && src) {
    uint32_t n_cb = 0;
    to_host(&n_cb, &csc->n_cblocks, sizeof(uint32_t));
    if (n_cb) {
      launch("contact_grid_to_reference");
      contact_grid_to_reference_kernel<T><<<blocks_for((size_t)n_cb * 64), kThreads, 0, stream>>>(n_cb, comps, cb_list,
                                                                                                 src, buf);
      GMPM_CUDA(cudaGetLastError());
    }
  }
  return buf;
}

}  // namespace gmpm
