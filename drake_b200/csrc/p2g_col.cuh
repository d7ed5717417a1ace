// p2g_col.cuh -- P2G, "column" formulation (replaces particle_to_grid_kernel :418-543 and clean_grid_kernel :585-602).
//
// Same contract as p2g_tile_kernel (kernels.cuh): one WARP per work record (a chunk of <= kChunk particles of one
// particle block, sorted by cell key), a private 6^3 x 4 shared-memory tile, no atomics, fixed summation order; the
// tile goes to tiles[t][4][216] and the grid kernel sums the <= 8 tiles over each grid block.
//
// What changed is who owns what in phase (b).  The tile kernel makes lanes = the 27 stencil nodes of ONE particle:
// every lane then consumes 64 bytes of operands per particle and the kernel is bound by the shared-memory pipe
// (27 wavefronts per particle, profiles/r01_v6_*: a shared load costs (bytes per lane)/4 wavefronts whatever the
// broadcast pattern).  Here a lane owns a (j,k) COLUMN of a particle's stencil and keeps its three i-nodes x 4 channels
// in registers; 9 lanes serve one particle, so the warp works on THREE particles at a time (3 sub-warps x 9 lanes) and
// a lane consumes 64 bytes per particle for 3 nodes instead of 1:
//     A'   = wy_j wz_k m                G_d = wy_j wz_k (a'_d + By_d j + Bz_d k)      (9 x 4 scalars, phase (a))
//     fx, bx_d = Bx_d / m                                                             (4 scalars per particle)
//     node i:  mass += wx_i A'          mom_d += wx_i (G_d + i A' bx_d)               wx_i recomputed from fx
// Each sub-warp walks its own contiguous third of the chunk, accumulates a run of particles of the same cell in
// registers and adds the run to the tile when the cell changes; the three sub-warps flush one after the other (their
// stencils overlap), 9 lanes x 12 read-modify-writes each.  With lanes = (j,k) the 9 columns fall into distinct banks
// of the UNPADDED [4][216] tile (word index 6j + k), so the tile stays the contiguous 6912-byte (fp64) image the grid
// kernel reads and can leave the SM as ONE bulk copy (cp.async.bulk.global.shared::cta) while the next chunk's
// phase (a) runs.
//
// Measured (profiles/r02_p2g_ab.txt, r02_p2g_col_ncu_summary.txt; 1M particles, fp64): shared-memory wavefronts per
// particle 27.0 -> 17.8 as designed, but the kernel does NOT get faster (157 us with 10 rows per sub-warp, 141 us with
// 7, against 145 us for the tile kernel; 101 vs 78 us at fp32): with 11-14 one-warp CTAs per SM neither kernel is bound
// by a pipe (shared-memory pipe 45 %, fp64 23 %, issue slots 38 %) but by dependent-issue and scoreboard latency at
// 16 % occupancy, and this formulation executes more instructions per particle (58 vs 49).  It stays as a switchable
// variant (GMPM_P2G=col|colbulk); the tile kernel is the default.
#pragma once
#include "kernels.cuh"

namespace gmpm {

#ifndef GMPM_COL_PER
#define GMPM_COL_PER 7
#endif
constexpr int kColPer = GMPM_COL_PER;   // particles per sub-warp and batch (3 * kColPer <= 32)
constexpr int kColRows = 3 * kColPer;   // staged particles per batch
#ifndef GMPM_COL_PREFETCH
#define GMPM_COL_PREFETCH 1
#endif
constexpr bool kColPrefetch = GMPM_COL_PREFETCH != 0;  // L2 prefetch of the next item's particle records

// Stage row: [9 x 16 B: (A', G0)] [9 x 16 B: (G1, G2)] [(fx, bx0) (bx1, bx2)] for double (two 16-byte planes: a
// 32-byte struct per lane would put lanes jk and jk + 4 into the same banks), [9 x 16 B: (A', G0, G1, G2)] [(fx, bx0,
// bx1, bx2)] for float.  Row stride = 16 (mod 128) bytes and the rows of the three sub-warps adjacent: the 27 loads
// of one instruction fall into distinct banks.
template <typename T> struct ColStage;
template <> struct ColStage<double> {
  static constexpr int kRowBytes = 400;
  __device__ static __forceinline__ void store_a(unsigned char* row, int jk, double a, double g0, double g1, double g2) {
    *reinterpret_cast<double2*>(row + jk * 16) = make_double2(a, g0);
    *reinterpret_cast<double2*>(row + 144 + jk * 16) = make_double2(g1, g2);
  }
  __device__ static __forceinline__ void store_z(unsigned char* row, double fx, double b0, double b1, double b2) {
    *reinterpret_cast<double2*>(row + 288) = make_double2(fx, b0);
    *reinterpret_cast<double2*>(row + 304) = make_double2(b1, b2);
  }
  __device__ static __forceinline__ void load(const unsigned char* row, int jk, double (&q)[8]) {
    const double2 p0 = *reinterpret_cast<const double2*>(row + jk * 16);
    const double2 p1 = *reinterpret_cast<const double2*>(row + 144 + jk * 16);
    const double2 z0 = *reinterpret_cast<const double2*>(row + 288);
    const double2 z1 = *reinterpret_cast<const double2*>(row + 304);
    q[0] = p0.x; q[1] = p0.y; q[2] = p1.x; q[3] = p1.y; q[4] = z0.x; q[5] = z0.y; q[6] = z1.x; q[7] = z1.y;
  }
};
template <> struct ColStage<float> {
  static constexpr int kRowBytes = 272;
  __device__ static __forceinline__ void store_a(unsigned char* row, int jk, float a, float g0, float g1, float g2) {
    *reinterpret_cast<float4*>(row + jk * 16) = make_float4(a, g0, g1, g2);
  }
  __device__ static __forceinline__ void store_z(unsigned char* row, float fx, float b0, float b1, float b2) {
    *reinterpret_cast<float4*>(row + 144) = make_float4(fx, b0, b1, b2);
  }
  __device__ static __forceinline__ void load(const unsigned char* row, int jk, float (&q)[8]) {
    const float4 p = *reinterpret_cast<const float4*>(row + jk * 16);
    const float4 z = *reinterpret_cast<const float4*>(row + 144);
    q[0] = p.x; q[1] = p.y; q[2] = p.z; q[3] = p.w; q[4] = z.x; q[5] = z.y; q[6] = z.z; q[7] = z.w;
  }
};

template <typename T>
struct P2GColSmem {
  T tile[4][kTileNodes];
  alignas(16) unsigned char stage[kColRows * ColStage<T>::kRowBytes];
};

// Phase (a) for one particle: the staged row of the column formulation.  Returns the reached-neighbour mask.
template <typename T>
__device__ __forceinline__ uint32_t p2g_derive_col(const P2GRaw<T>& r, unsigned char* row, T dt, const Consts<T>& k) {
  uint32_t b[3];
  T fx[3], w[3][3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    b[d] = base_node(r.p[d], k.dx_inv, k.domain_size);
    fx[d] = r.p[d] * k.dx_inv - T(b[d]);
    w[0][d] = T(0.5) * (T(1.5) - fx[d]) * (T(1.5) - fx[d]);
    w[1][d] = T(0.75) - (fx[d] - T(1)) * (fx[d] - T(1));
    w[2][d] = T(0.5) * (fx[d] - T(0.5)) * (fx[d] - T(0.5));
  }
  const T mass = r.vol * k.density;
  const T inv_mass = mass > T(0) ? T(1) / mass : T(0);
  T B[9];
#pragma unroll
  for (int e = 0; e < 9; ++e) B[e] = r.C[e] * mass;
  T a[3] = {mass * r.v[0], mass * r.v[1], mass * r.v[2]};
#pragma unroll
  for (int d = 0; d < 3; ++d)
    if (d == k.gravity_axis) a[d] += mass * k.gravity * dt;
  if (r.solid) {
    const T cs = -dt * k.d_inv;
    B[0] += cs * r.ex[0]; B[4] += cs * r.ex[1]; B[8] += cs * r.ex[2];
    B[1] += cs * r.ex[3]; B[3] += cs * r.ex[3];
    B[2] += cs * r.ex[4]; B[6] += cs * r.ex[4];
    B[5] += cs * r.ex[5]; B[7] += cs * r.ex[5];
  } else if (r.face) {
    const T cs = -dt * k.d_inv;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) B[i * 3 + j] += cs * (r.ex[i] * r.ex[3 + j]);
  } else {
#pragma unroll
    for (int d = 0; d < 3; ++d) a[d] += r.ex[d] * dt;
  }
  T Bx[3], By[3], Bz[3], al[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    Bx[d] = B[d * 3 + 0] * k.dx;
    By[d] = B[d * 3 + 1] * k.dx;
    Bz[d] = B[d * 3 + 2] * k.dx;
    al[d] = a[d] - (Bx[d] * fx[0] + By[d] * fx[1] + Bz[d] * fx[2]);
  }
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      const T A = w[j][1] * w[kk][2];
      ColStage<T>::store_a(row, j * 3 + kk, A * mass, A * (al[0] + By[0] * T(j) + Bz[0] * T(kk)),
                           A * (al[1] + By[1] * T(j) + Bz[1] * T(kk)), A * (al[2] + By[2] * T(j) + Bz[2] * T(kk)));
    }
  ColStage<T>::store_z(row, fx[0], Bx[0] * inv_mass, Bx[1] * inv_mass, Bx[2] * inv_mass);
  uint32_t reach = 1u;  // bit (dx*4+dy*2+dz) per reached neighbour block
  if ((b[0] & 3u) >= 2u) reach |= reach << 4;
  if ((b[1] & 3u) >= 2u) reach |= reach << 2;
  if ((b[2] & 3u) >= 2u) reach |= reach << 1;
  // a massless particle (an inactive spare of the migration, csrc/migrate.cuh) deposits zeros: it must not make a
  // grid block "touched" -- the exchange plans the reachable blocks from the mass on the grid
  return mass > T(0) ? reach : 0u;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <typename T, bool BULK>
__global__ void __launch_bounds__(kP2GThreads)
p2g_col_kernel(uint32_t N, uint2 n_faces, const uint32_t* __restrict__ order,
               const uint32_t* __restrict__ sorted_key, const WorkRec* __restrict__ pb_work, uint32_t work_cap,
               const T* __restrict__ pos, const T* __restrict__ vel, const T* __restrict__ vol,
               const T* __restrict__ C, const int* __restrict__ pid, const T* __restrict__ forces,
               const T* __restrict__ taus, T* __restrict__ tiles, uint32_t* __restrict__ touched_epoch,
               uint32_t* __restrict__ touched_list, uint32_t epoch, DevCounters* ctr, T dt, Consts<T> k,
               uint32_t* __restrict__ nbr_list = nullptr) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  P2GColSmem<T>& sm = *reinterpret_cast<P2GColSmem<T>*>(smem_raw);
  constexpr int kRow = ColStage<T>::kRowBytes;
  const int lane = threadIdx.x;
  T* tile = &sm.tile[0][0];
  const uint32_t n_big = ctr->n_big, n_work = n_big + ctr->n_small;
  if (n_work == 0) return;
  // phase (b) role: sub-warp g, column (nj, nk)
  const bool col_lane = lane < 27;
  const int g = col_lane ? lane / 9 : 0;
  const int jk = col_lane ? lane % 9 : 0, nj = jk / 3, nk = jk % 3;
  // phase (a) role: particle pr of sub-warp pg in the batch
  const bool part_lane = lane < kColRows;
  const uint32_t pg = part_lane ? lane / kColPer : 0u, pr = lane % kColPer;
  unsigned char* my_row = sm.stage + (pr * 3 + pg) * kRow;
  const uint32_t blocks_per_axis = (uint32_t)k.domain_size >> 2;

  constexpr uint32_t kNone = 0xFFFFFFFFu;
  using Item = WorkRec;
  auto ticket = [&]() -> uint32_t {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(&ctr->work_p2g, 1u);
    return t;
  };
  auto resolve = [&](uint32_t t_lane0) -> Item {
    const uint32_t t = __shfl_sync(kFullMask, t_lane0, 0);
    if (t >= n_work) return Item{kNone, 0u, 0u, 0u};
    return pb_work[t < n_big ? t : work_cap - 1u - (t - n_big)];
  };
  // sub-warp pg walks the third [pg L, (pg + 1) L) of the chunk, L = ceil(n / 3), kColPer particles per batch
  auto n_batches = [&](const Item& it) -> uint32_t {
    return ((it.end - it.begin + 2u) / 3u + kColPer - 1u) / kColPer;
  };
  auto slot_of = [&](const Item& it, uint32_t b, uint32_t& s, uint32_t& key) -> bool {
    if (it.tile == kNone || !part_lane) return false;
    const uint32_t n = it.end - it.begin, L = (n + 2u) / 3u;
    const uint32_t off = b * kColPer + pr, p = pg * L + off;
    if (off >= L || p >= n) return false;
    s = order[it.begin + p];
    key = sorted_key[it.begin + p];
    return true;
  };

  Item cur = resolve(ticket());
  if (cur.tile == kNone) return;
  Item nxt = resolve(ticket());
  uint32_t nb_cur = n_batches(cur);
  // `raw` holds the records of the batch being processed, (v_b, s_b, key_b) the indices of the one after
  uint32_t s_a = 0, key_a = 0, s_b = 0, key_b = 0;
  bool v_raw = slot_of(cur, 0, s_a, key_a);
  bool v_b = 1u < nb_cur ? slot_of(cur, 1, s_b, key_b) : slot_of(nxt, 0, s_b, key_b);
  P2GRaw<T> raw;
  p2g_load(raw, v_raw, s_a, N, n_faces, pos, vel, vol, C, pid, forces, taus);
  uint32_t key_cur = key_a;

  T acc[3][4];
  // records of the item after the current one -> L2 (the physical particle order is the key order of the previous
  // substep, so the records of sorted positions [begin, end) sit around slots [begin, end)): the batch loads then
  // cost an L2 round trip instead of an HBM one
  auto prefetch_item = [&](const Item& it) {
    if (it.tile == kNone) return;
    constexpr uint32_t kPer = 128u / sizeof(T);  // scalars per 128-byte line
    const uint32_t lo = it.begin & ~(kPer - 1u), lines = (it.end - lo + kPer - 1u) / kPer;
    if ((uint32_t)lane < lines) {
      const size_t s = min((size_t)lo + (size_t)lane * kPer, (size_t)N - 1u);
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pos + (size_t)d * N + s));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(vel + (size_t)d * N + s));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(forces + (size_t)d * N + s));
      }
#pragma unroll
      for (int e = 0; e < 9; ++e) asm volatile("prefetch.global.L2 [%0];" ::"l"(C + (size_t)e * N + s));
#pragma unroll
      for (int e = 0; e < 6; ++e) asm volatile("prefetch.global.L2 [%0];" ::"l"(taus + (size_t)e * N + s));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(vol + s));
    }
  };
  if (!BULK) {
    for (int i = lane; i < 4 * kTileNodes; i += 32) tile[i] = T(0);
  }
  for (;;) {
    const uint32_t blk = cur.block;
    uint32_t reach = 0;
    int run_cell = -1, carry = -1;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[i][c] = T(0);
    const uint32_t t_pending = ticket();  // the item after `nxt`
    Item nn{kNone, 0u, 0u, 0u};
    if (kColPrefetch) prefetch_item(nxt);
    __syncwarp();

    auto flush = [&]() {
      const int base = ((run_cell >> 4) & 3) * 36 + (((run_cell >> 2) & 3) + nj) * 6 + ((run_cell & 3) + nk);
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int c = 0; c < 4; ++c) tile[c * kTileNodes + base + 36 * i] += acc[i][c];
    };

    for (uint32_t b = 0; b < nb_cur; ++b) {
      const bool last_batch = b + 1u >= nb_cur;
      // ---- phase (a): lanes = particles
      const int my_cell = (int)(key_cur & 63u);
      if (v_raw) reach |= p2g_derive_col(raw, my_row, dt, k);
      int prev = __shfl_up_sync(kFullMask, my_cell, 1);
      if (pr == 0) prev = carry;
      const uint32_t heads = __ballot_sync(kFullMask, v_raw && my_cell != prev);
      const uint32_t vmask = __ballot_sync(kFullMask, v_raw);
      carry = __shfl_sync(kFullMask, my_cell, (int)(pg * kColPer + kColPer - 1u));
      if (BULK && b == 0) {
        // the previous tile is still being read by the bulk copy engine: wait for that, then clear it
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        for (int i = lane; i < 4 * kTileNodes; i += 32) tile[i] = T(0);
      }
      __syncwarp();
      // ---- prefetch: records of the next batch, indices of the one after
      if (last_batch && nxt.tile != kNone) nn = resolve(t_pending);
      v_raw = v_b;
      p2g_load(raw, v_raw, s_b, N, n_faces, pos, vel, vol, C, pid, forces, taus);
      key_cur = key_b;
      if (b + 2u < nb_cur) v_b = slot_of(cur, b + 2u, s_b, key_b);
      else if (!last_batch) v_b = slot_of(nxt, 0, s_b, key_b);
      else if (nxt.tile != kNone && 1u < n_batches(nxt)) v_b = slot_of(nxt, 1, s_b, key_b);
      else v_b = slot_of(nn, 0, s_b, key_b);
      // ---- phase (b): lanes = (sub-warp, column).  Two operand sets in flight (ping-pong): the loads of step r + 1
      // are issued before the math of step r.
      const int rows0 = __popc(vmask & ((1u << kColPer) - 1u));  // sub-warp 0 always holds the longest third
      const uint32_t my_heads = col_lane ? (heads >> (g * kColPer)) & ((1u << kColPer) - 1u) : 0u;
      const uint32_t my_valid = col_lane ? (vmask >> (g * kColPer)) & ((1u << kColPer) - 1u) : 0u;
      const uint32_t any_heads = (heads | (heads >> kColPer) | (heads >> (2 * kColPer))) & ((1u << kColPer) - 1u);
      const unsigned char* rowp = sm.stage + g * kRow;
      auto step = [&](const T (&q)[8], int r) {
        if ((any_heads >> r) & 1u) {  // warp-uniform: some sub-warp starts a new run at this step
          const int newcell = __shfl_sync(kFullMask, my_cell, g * kColPer + r);
#pragma unroll
          for (int s = 0; s < 3; ++s)
            if ((heads >> (kColPer * s + r)) & 1u) {  // the sub-warps flush one after the other
              if (col_lane && g == s && run_cell >= 0) flush();
              __syncwarp();
            }
          if ((my_heads >> r) & 1u) {
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
              for (int c = 0; c < 4; ++c) acc[i][c] = T(0);
            run_cell = newcell;
          }
        }
        if ((my_valid >> r) & 1u) {
          const T fx = q[4];
          const T w0 = T(0.5) * (T(1.5) - fx) * (T(1.5) - fx);
          const T w1 = T(0.75) - (fx - T(1)) * (fx - T(1));
          const T w2 = T(0.5) * (fx - T(0.5)) * (fx - T(0.5));
          acc[0][0] += w0 * q[0];
          acc[1][0] += w1 * q[0];
          acc[2][0] += w2 * q[0];
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            const T u = q[0] * q[5 + d];
            acc[0][1 + d] += w0 * q[1 + d];
            acc[1][1 + d] += w1 * (q[1 + d] + u);
            acc[2][1 + d] += w2 * (q[1 + d] + T(2) * u);
          }
        }
      };
      T qa[8], qb[8];
      ColStage<T>::load(rowp, jk, qa);
      for (int r = 0; r < rows0; r += 2) {
        ColStage<T>::load(rowp + min(r + 1, rows0 - 1) * 3 * kRow, jk, qb);
        step(qa, r);
        if (r + 1 < rows0) {
          ColStage<T>::load(rowp + min(r + 2, rows0 - 1) * 3 * kRow, jk, qa);
          step(qb, r + 1);
        }
      }
      __syncwarp();
    }
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      if (col_lane && g == s && run_cell >= 0) flush();
      __syncwarp();
    }
    // ---- the tile leaves the SM (one bulk copy, or coalesced stores) and the touched grid blocks are published
    T* out = tiles + (size_t)cur.tile * (4 * kTileNodes);
    if (BULK) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the copy engine
      __syncwarp();
      if (lane == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out), "r"(smem_u32(tile)),
                     "r"((uint32_t)(4 * kTileNodes * sizeof(T)))
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    } else {
      for (int i = lane; i < 4 * kTileNodes; i += 32) {  // write the tile out and leave it cleared for the next item
        out[i] = tile[i];
        tile[i] = T(0);
      }
    }
    reach = __reduce_or_sync(kFullMask, reach);
    if (lane < 8 && ((reach >> lane) & 1u)) {
      uint32_t bx, by, bz;
      block_coords(blk, bx, by, bz);
      bx += (lane >> 2) & 1u;
      by += (lane >> 1) & 1u;
      bz += lane & 1u;
      if (bx < blocks_per_axis && by < blocks_per_axis && bz < blocks_per_axis) {
        const uint32_t gb = morton3(bx, by, bz);
        if (atomicExch(&touched_epoch[gb], epoch) != epoch) {
          const uint32_t idx = atomicAdd(&ctr->n_touched, 1u);
          touched_list[idx] = gb;
          if (nbr_list) nbr_list[idx] = 0u;  // filled by the grid assembly (multi-GPU planning)
        }
      }
    }
    __syncwarp();
    if (nxt.tile == kNone) break;
    cur = nxt;
    nxt = nn;
    nb_cur = n_batches(cur);
  }
  if (BULK) {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the last tile has landed
    __syncwarp();
  }
}

}  // namespace gmpm
