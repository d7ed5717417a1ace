// kernels.cuh -- the per-timestep MPM kernels of libgmpm_b200 (sm_100a).
//
// Data layout (DESIGN.md "Data layout in HBM"):
//   particles : planar SoA by PHYSICAL slot: pos[3][N] vel[3][N] C[9][N] vol[N] pid[N], double
//               buffered; forces[3][N], taus[6][N] (the two factors of the rank-one stress) by slot.  slot_of_pid[N] is the inverse of pid[].
//   faces     : planar by face id: tri[3][F] (original particle ids), F[9][F], Dm_inv[4][F].
//   binning   : key[N] (block-major Morton cell key of every slot), order[N] = slots sorted by key,
//               per-particle-block ranges pb_start/pb_end[n_blocks], compact list pb_list[].
//   grid      : dense, block-major, channel-planar inside each 4^3 block:
//               g_mass[b*64+l], g_vel[(b*3+d)*64+l] (momentum after P2G, velocity after update),
//               g_vstar likewise.  P2G tiles: tiles[t][4][216] (6^3 nodes of particle block t).
//
// Every kernel cites the reference kernel (multibody/gpu_mpm/cuda_mpm_kernels.cuh) it replaces.
#pragma once
#include "common.cuh"
#include "math.cuh"
#include "sort.cuh"
#include "halo_window.cuh"

namespace gmpm {

constexpr int kThreads = 256;
#ifndef GMPM_CHUNK
#define GMPM_CHUNK 128
#endif
constexpr int kChunk = GMPM_CHUNK;  // particles per unit of work (a particle block is split into chunks of <= kChunk)

// Device-resident counters (no host round trips inside a substep).
struct DevCounters {
  uint32_t n_pblocks;      // particle blocks of the current binning
  uint32_t n_touched;      // touched grid blocks of the current P2G
  uint32_t n_touched_prev; // ... of the previous P2G (for stale-block cleaning)
  uint32_t work_p2g;       // dynamic work queues of the persistent kernels
  uint32_t work_g2p;
  uint32_t work_grid;
  uint32_t n_big;          // work records of blocks with more particles than the average (front of pb_work)
  uint32_t n_small;        // ... the others (back of pb_work, growing down)
  uint32_t n_tiles;        // P2G tiles handed out (one per work record)
  uint32_t win_lo[3];      // bounding box of the particle blocks of the LAST binning (dilated Morton fields), from
  uint32_t win_hi[3];      // which the next binning predicts its sort window; lo = 0xFFFFFFFF: none yet
  uint32_t win_valid;      // win_lo / win_hi hold the box of a finished binning (written only by its last phase)
  uint32_t win_violation;  // a key fell outside the predicted window (the sort fell back to absolute keys)
  uint32_t win_fallbacks;  // how often that happened (instrumentation)
  uint32_t sort_passes;    // passes the last binning sort ran (instrumentation)
  uint32_t phase_ns[8];    // binning kernel: ns from start to the end of window / sort / ranges / records (block 0)
};

// One unit of work of the per-particle-block kernels: tile index, particle range, grid block id.
struct alignas(16) WorkRec {
  uint32_t tile, begin, end, block;
};

// Rolls the touched-block list and resets the P2G work queue (device side, no host round trip).
__global__ void p2g_prologue_kernel(DevCounters* ctr) {
  ctr->n_touched_prev = ctr->n_touched;
  ctr->n_touched = 0;
  ctr->work_p2g = 0;
  ctr->work_g2p = 0;
  ctr->work_grid = 0;
}


// ------------------------------------------------------------------------------------------------
// Setup: rest shape, F = Q, volumes (replaces initialize_fem_state_kernel, :13-70).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void init_fem_state_kernel(uint32_t n_faces, uint32_t N, const int* __restrict__ tri, T* pos, T* vel,
                                      T* vol, T* Dm_inv, T* Fd, T dx) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_faces) return;
  const int v0 = tri[f], v1 = tri[n_faces + f], v2 = tri[2 * n_faces + f];  // slots == pids at setup
  T x0[3], x1[3], x2[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    x0[d] = pos[d * N + v0];
    x1[d] = pos[d * N + v1];
    x2[d] = pos[d * N + v2];
    pos[d * N + f] = (x0[d] + x1[d] + x2[d]) / T(3);
    vel[d * N + f] = (vel[d * N + v0] + vel[d * N + v1] + vel[d * N + v2]) / T(3);
  }
  T D0[3], D1[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    D0[d] = x1[d] - x0[d];
    D1[d] = x2[d] - x0[d];
  }
  const T Ds[6] = {D0[0], D1[0], D0[1], D1[1], D0[2], D1[2]};
  T Q[9], R[6];
  qr3<T, 2>(Ds, Q, R);
  const T Dm[4] = {R[0], R[1], T(0), R[3]};
  T Dmi[4];
  inverse2(Dm, Dmi);
#pragma unroll
  for (int e = 0; e < 4; ++e) Dm_inv[e * n_faces + f] = Dmi[e];
#pragma unroll
  for (int e = 0; e < 9; ++e) Fd[e * n_faces + f] = Q[e];
  const T cx = D0[1] * D1[2] - D1[1] * D0[2];
  const T cy = D0[2] * D1[0] - D1[2] * D0[0];
  const T cz = D0[0] * D1[1] - D1[0] * D0[1];
  const T v4 = tsqrt(cx * cx + cy * cy + cz * cz) / T(8) * dx;
  atomicAdd(&vol[f], v4);
  atomicAdd(&vol[v0], v4);
  atomicAdd(&vol[v1], v4);
  atomicAdd(&vol[v2], v4);
}

// ------------------------------------------------------------------------------------------------
// Keys (replaces compute_base_cell_node_index_kernel, :365-382).  One key per physical slot.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void compute_keys_kernel(uint32_t N, const T* __restrict__ pos, uint32_t* __restrict__ key,
                                    Consts<T> k) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const uint32_t bx = base_node(pos[i], k.dx_inv, k.domain_size);
  const uint32_t by = base_node(pos[N + i], k.dx_inv, k.domain_size);
  const uint32_t bz = base_node(pos[2 * N + i], k.dx_inv, k.domain_size);
  key[i] = cell_key(bx, by, bz);
}

// Visible-order keys for the reference-compatible sort bookkeeping: vis_key[v] = key[slot(v)].
__global__ void gather_visible_keys_kernel(uint32_t N, const uint32_t* __restrict__ key,
                                           const int* __restrict__ pid_of_vis, const int* __restrict__ slot_of_pid,
                                           uint32_t* __restrict__ vis_key, uint32_t* __restrict__ vis_id) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  vis_key[v] = key[slot_of_pid[pid_of_vis[v]]];
  vis_id[v] = v;
}

// After the reference-compatible stable sort: permute the visible order
// (the integer part of compute_sorted_state_kernel, :384-416: pids and index_mappings).
__global__ void permute_visible_kernel(uint32_t N, const uint32_t* __restrict__ sort_ids,
                                       const int* __restrict__ pid_of_vis_old, int* __restrict__ pid_of_vis_new,
                                       int* __restrict__ vis_of_pid) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const int p = pid_of_vis_old[sort_ids[v]];
  pid_of_vis_new[v] = p;
  vis_of_pid[p] = (int)v;
}

// ------------------------------------------------------------------------------------------------
// Binning
// ------------------------------------------------------------------------------------------------
// The whole particle binning in one cooperative launch: stable radix sort of (cell key, slot) on all
// key bits, then the per-block ranges / block list (build_pblocks) and the chunked work records
// (classify_pblocks), with grid barriers in between.  `sorted` = index of the result buffer pair.
struct BinArgs {
  SortArgs sort;
  uint32_t N, epoch, work_cap;
  int roll_touched;  // binning on behalf of P2G: also do p2g_prologue_kernel's job (saves a launch)
  int window;        // sort keys relative to the particles' bounding box (grids whose keys need > 2 passes)
  int key_bits;      // bits of an absolute cell key
  uint32_t *pb_start, *pb_end, *pb_epoch, *pb_list;
  uint2* pb_tile;
  WorkRec* pb_work;
  DevCounters* ctr;
};

__device__ __forceinline__ uint64_t global_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(kSortThreads, 2) bin_particles_kernel(BinArgs a) {
  extern __shared__ __align__(16) uint32_t sort_smem[];
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const bool stamp = blockIdx.x == 0 && threadIdx.x == 0;
  const uint64_t t_begin = stamp ? global_ns() : 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    a.ctr->n_pblocks = 0;
    a.ctr->n_big = 0;
    a.ctr->n_small = 0;
    a.ctr->n_tiles = 0;
    if (a.roll_touched) {
      a.ctr->n_touched_prev = a.ctr->n_touched;
      a.ctr->n_touched = 0;
      a.ctr->work_p2g = 0;
      a.ctr->work_g2p = 0;
      a.ctr->work_grid = 0;
    }
  }
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  SortWindow win{};
  if (a.window) {
    // Sort window = bounding box of the particle blocks, per axis on the DILATED fields of the Morton block id
    // (dilation is monotonic, so min / max need no decoding).  Normally it is PREDICTED: the box of the previous
    // binning (left in ctr->win_* by its last phase) grown by one block -- a particle moves less than a cell per
    // substep -- which costs no pass over the keys and no barrier; pass 0 of the sort verifies it and falls back to
    // the absolute keys if a key lies outside.  Only the first binning of a state measures the box up front.
    const int axis_bits = (a.key_bits - 6) / 3;
    const uint32_t axis_max = (1u << axis_bits) - 1u;
    uint32_t lo_i[3], hi_i[3];
    // win_valid is written by the last phase of a launch only: every thread of this one reads the same value
    // (win_lo / win_hi themselves are accumulators that the measuring path below updates while CTAs still start)
    const bool predicted = __ldcg(&a.ctr->win_valid) != 0u;
    if (predicted) {
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const uint32_t l = contract_bits(__ldcg(&a.ctr->win_lo[d]) >> (2 - d));
        const uint32_t h = contract_bits(__ldcg(&a.ctr->win_hi[d]) >> (2 - d));
        lo_i[d] = l > 0u ? l - 1u : 0u;
        hi_i[d] = h < axis_max ? h + 1u : axis_max;
      }
    } else {
      uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
      for (uint32_t i = tid; i < a.N; i += nth) {
        const uint32_t b = __ldcg(a.sort.keys_in + i) >> 6;
        const uint32_t c[3] = {b & kMortonX, b & kMortonY, b & kMortonZ};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          lo[d] = min(lo[d], c[d]);
          hi[d] = max(hi[d], c[d]);
        }
      }
      __shared__ uint32_t s_lo[3], s_hi[3];
      if (threadIdx.x < 3) {
        s_lo[threadIdx.x] = 0xFFFFFFFFu;
        s_hi[threadIdx.x] = 0u;
      }
      __syncthreads();
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        lo[d] = __reduce_min_sync(kFullMask, lo[d]);
        hi[d] = __reduce_max_sync(kFullMask, hi[d]);
        if ((threadIdx.x & 31) == 0) {
          atomicMin(&s_lo[d], lo[d]);
          atomicMax(&s_hi[d], hi[d]);
        }
      }
      __syncthreads();
      if (threadIdx.x < 3) {  // one pair of global atomics per CTA and axis
        atomicMin(&a.ctr->win_lo[threadIdx.x], s_lo[threadIdx.x]);
        atomicMax(&a.ctr->win_hi[threadIdx.x], s_hi[threadIdx.x]);
      }
      grid.sync();
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        lo_i[d] = contract_bits(__ldcg(&a.ctr->win_lo[d]) >> (2 - d));
        hi_i[d] = contract_bits(__ldcg(&a.ctr->win_hi[d]) >> (2 - d));
      }
    }
    uint32_t ext = 1;
#pragma unroll
    for (int d = 0; d < 3; ++d) ext = max(ext, hi_i[d] - lo_i[d] + 1u);
    int w = 0;
    while ((1u << w) < ext) ++w;
    const int rel_bits = 3 * w + 6;
    if (rel_bits < a.key_bits) {
      win.on = true;
      win.ox = morton3(lo_i[0], 0u, 0u);
      win.oy = morton3(0u, lo_i[1], 0u);
      win.oz = morton3(0u, 0u, lo_i[2]);
      win.n_passes = plan_sort_passes(rel_bits, win.pass);
      win.final_buf = (a.sort.n_passes - 1) & 1;
      win.rel_bits = rel_bits;
      win.violation = predicted ? &a.ctr->win_violation : nullptr;
      if (win.n_passes >= a.sort.n_passes) win.on = false;  // no pass saved: keep the absolute keys
    }
  }
  if (stamp) a.ctr->phase_ns[0] = (uint32_t)(global_ns() - t_begin);
  win.stamp_out = stamp ? &a.ctr->phase_ns[4] : nullptr;
  win.t0 = t_begin;
  const int r = radix_sort_device(a.sort, sort_smem, grid, &win);  // ends with a grid barrier
  if (stamp) a.ctr->phase_ns[1] = (uint32_t)(global_ns() - t_begin);
  const uint32_t* sorted_key = a.sort.keys[r];
  if (blockIdx.x == 0 && threadIdx.x == 0) a.ctr->sort_passes = win.on ? win.n_passes : a.sort.n_passes;
  if (a.window && blockIdx.x == 0 && threadIdx.x < 3) {  // everyone has read the window: reset the accumulators
    a.ctr->win_lo[threadIdx.x] = 0xFFFFFFFFu;            // (the last phase below leaves the new box in them)
    a.ctr->win_hi[threadIdx.x] = 0u;
    if (threadIdx.x == 0 && a.ctr->win_violation) {
      a.ctr->win_violation = 0u;
      ++a.ctr->win_fallbacks;
    }
  }
  // ---- block ranges + list of particle blocks (uniform trip count: the warp-aggregated append needs all lanes)
  for (uint32_t i0 = 0; i0 < a.N; i0 += nth) {
    const uint32_t i = i0 + tid;
    uint32_t b = 0;
    bool head = false;
    if (i < a.N) {
      b = sorted_key[i] >> 6;
      head = (i == 0) || ((sorted_key[i - 1] >> 6) != b);
      const bool tail = (i + 1 == a.N) || ((sorted_key[i + 1] >> 6) != b);
      if (win.on && (head || tail)) b = window_block_abs(b, win);  // sorted keys are window-relative
      if (head) a.pb_start[b] = i;
      if (tail) a.pb_end[b] = i + 1;
    }
    const uint32_t m = __ballot_sync(kFullMask, head);
    if (m) {
      const int lane = threadIdx.x & 31;
      uint32_t base = 0;
      if (lane == (__ffs(m) - 1)) base = atomicAdd(&a.ctr->n_pblocks, (uint32_t)__popc(m));
      base = __shfl_sync(kFullMask, base, __ffs(m) - 1);
      if (head) {
        a.pb_list[base + __popc(m & ((1u << lane) - 1u))] = b;
        a.pb_epoch[b] = a.epoch;
      }
    }
  }
  grid.sync();
  if (stamp) a.ctr->phase_ns[2] = (uint32_t)(global_ns() - t_begin);
  // ---- chunked work records (+ the bounding box of the particle blocks, for the next launch's window)
  const uint32_t n_pb = a.ctr->n_pblocks;
  if (a.window) {
    uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
    for (uint32_t t = tid; t < n_pb; t += nth) {
      const uint32_t b = a.pb_list[t];
      const uint32_t c[3] = {b & kMortonX, b & kMortonY, b & kMortonZ};
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        lo[d] = min(lo[d], c[d]);
        hi[d] = max(hi[d], c[d]);
      }
    }
    if (tid == 0) a.ctr->win_valid = n_pb ? 1u : 0u;
    if (tid - (threadIdx.x & 31) < n_pb) {  // warps that saw a block
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        lo[d] = __reduce_min_sync(kFullMask, lo[d]);
        hi[d] = __reduce_max_sync(kFullMask, hi[d]);
        if ((threadIdx.x & 31) == 0) {
          atomicMin(&a.ctr->win_lo[d], lo[d]);
          atomicMax(&a.ctr->win_hi[d], hi[d]);
        }
      }
    }
  }
  for (uint32_t t = tid; t < n_pb; t += nth) {
    const uint32_t blk = a.pb_list[t];
    const uint32_t b = a.pb_start[blk], e = a.pb_end[blk];
    const uint32_t n_full = (e - b) / kChunk, rest = (e - b) % kChunk;
    const uint32_t n_chunks = n_full + (rest ? 1u : 0u);
    const uint32_t tile0 = atomicAdd(&a.ctr->n_tiles, n_chunks);
    a.pb_tile[blk] = make_uint2(tile0, n_chunks);
    if (n_full) {
      const uint32_t w0 = atomicAdd(&a.ctr->n_big, n_full);
      for (uint32_t c = 0; c < n_full; ++c)
        a.pb_work[w0 + c] = WorkRec{tile0 + c, b + c * kChunk, b + (c + 1) * kChunk, blk};
    }
    if (rest)
      a.pb_work[a.work_cap - 1u - atomicAdd(&a.ctr->n_small, 1u)] = WorkRec{tile0 + n_full, b + n_full * kChunk, e, blk};
  }
  if (stamp) a.ctr->phase_ns[3] = (uint32_t)(global_ns() - t_begin);
}

// ------------------------------------------------------------------------------------------------
// Constitutive update (replaces calc_fem_state_and_force_kernel, :183-294).  One thread per face.
// ------------------------------------------------------------------------------------------------
// Face records (tri, F, Dm^-1) are kept in the SPATIAL order of their face particles (refreshed every
// few dozen substeps, State::refresh_face_order): record j belongs to face fid[j], consecutive
// records sit in consecutive particle slots and their vertices in nearby ones, so the slot-indexed
// loads, stores and atomics of a warp fall into a few 128-byte lines instead of ~16 each.
// Memory-latency design: the face record needs three dependent levels of loads (face -> particle ids ->
// slots -> vertex state).  The kernel is persistent (grid-stride over faces) and software-pipelined:
// the slot indices of the NEXT face of a thread are resolved while the current face is computed, and
// every load of the current face (F, Dm^-1, C, volume, the three vertices' x and v) is issued before
// the first use, so a face costs about one memory round trip instead of a dozen serialized ones.
// Cell keys of the face particles at the centroids the constitutive kernel is about to move them to (the same
// expression as fem_kernel below): with the keys known up front the particle binning -- latency bound, a fifth of
// the SMs' warps -- runs on a second stream UNDER the constitutive kernel instead of after it (State::calc_fem).
template <typename T>
__global__ void face_keys_kernel(uint32_t n_faces, uint32_t N, const int* __restrict__ fid, const int* __restrict__ tri,
                                 const int* __restrict__ slot_of_pid, const T* __restrict__ pos,
                                 uint32_t* __restrict__ key, Consts<T> k) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_faces) return;
  const int sf = slot_of_pid[fid[j]];
  const int s0 = slot_of_pid[tri[j]], s1 = slot_of_pid[tri[n_faces + j]], s2 = slot_of_pid[tri[2 * n_faces + j]];
  T xc[3];
#pragma unroll
  for (int d = 0; d < 3; ++d)
    xc[d] = (pos[(size_t)d * N + s0] + pos[(size_t)d * N + s1] + pos[(size_t)d * N + s2]) / T(3);
  key[sf] = cell_key(base_node(xc[0], k.dx_inv, k.domain_size), base_node(xc[1], k.dx_inv, k.domain_size),
                     base_node(xc[2], k.dx_inv, k.domain_size));
}

#ifndef GMPM_FEM_MIN_CTAS
#define GMPM_FEM_MIN_CTAS 3
#endif
constexpr int kFemThreads = 128;

struct FemIdx {
  int sf, s0, s1, s2;
};

template <typename T>
__global__ void __launch_bounds__(kFemThreads, GMPM_FEM_MIN_CTAS)
fem_kernel(uint32_t n_faces, uint32_t N, const int* __restrict__ fid, const int* __restrict__ tri,
           const int* __restrict__ slot_of_pid,
           const T* __restrict__ vol, const T* __restrict__ C, const T* __restrict__ Dm_inv, T* __restrict__ pos,
           T* __restrict__ vel, T* __restrict__ Fd, T* __restrict__ forces, T* __restrict__ taus,
           uint32_t* __restrict__ key, T dt, Consts<T> k, bool reverse) {
  const uint32_t stride = gridDim.x * blockDim.x;
  uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  // Traversal order: G2P, which ran just before, wrote the particle state in ascending slot order and the face
  // records are in slot order too, so walking them from the END meets the lines G2P wrote last while they are
  // still in L2 (walking from the start meets only lines the rest of G2P has already evicted).
  auto rec = [&](uint32_t ff) -> uint32_t { return reverse ? n_faces - 1u - ff : ff; };
  // two index levels, each resolved one iteration ahead of its use (an in-order warp stalls at the first
  // instruction whose operand is still in flight, so a dependent pair of loads must not share an iteration)
  auto level1 = [&](uint32_t ff) -> FemIdx {  // particle ids of the record: face particle, three vertices
    FemIdx r{0, 0, 0, 0};
    if (ff < n_faces) {
      const uint32_t j = rec(ff);
      r.sf = fid[j];
      r.s0 = tri[j];
      r.s1 = tri[n_faces + j];
      r.s2 = tri[2 * n_faces + j];
    }
    return r;
  };
  auto level2 = [&](const FemIdx& ids) -> FemIdx {  // their current slots
    return FemIdx{slot_of_pid[ids.sf], slot_of_pid[ids.s0], slot_of_pid[ids.s1], slot_of_pid[ids.s2]};
  };
  FemIdx cur = level2(level1(f));     // slots of face f
  FemIdx ids_n = level1(f + stride);  // particle ids of face f + stride
  for (; f < n_faces; f += stride) {
    const int sf = cur.sf, s0 = cur.s0, s1 = cur.s1, s2 = cur.s2;
    const uint32_t j = rec(f);  // record index
    // ---- every load of this face, before any use
    T x0[3], x1[3], x2[3], v0[3], v1[3], v2[3], F[9], Cm[9], Dmi[4];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      x0[d] = pos[(size_t)d * N + s0];
      x1[d] = pos[(size_t)d * N + s1];
      x2[d] = pos[(size_t)d * N + s2];
      v0[d] = vel[(size_t)d * N + s0];
      v1[d] = vel[(size_t)d * N + s1];
      v2[d] = vel[(size_t)d * N + s2];
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) {
      F[e] = Fd[(size_t)e * n_faces + j];
      Cm[e] = C[(size_t)e * N + sf];
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) Dmi[e] = Dm_inv[(size_t)e * n_faces + j];
    const T volume = vol[sf];
    // ---- indices of this thread's next faces (in flight during the math below)
    cur = level2(ids_n);                 // ids_n arrived during the previous iteration
    ids_n = level1(f + 2 * stride);

    T xc[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      xc[d] = (x0[d] + x1[d] + x2[d]) / T(3);
      pos[(size_t)d * N + sf] = xc[d];
      vel[(size_t)d * N + sf] = (v0[d] + v1[d] + v2[d]) / T(3);
    }
    // the face particle moved: refresh its cell key (unless face_keys_kernel has done so and the binning is
    // already running beside this kernel)
    if (key) key[sf] = cell_key(base_node(xc[0], k.dx_inv, k.domain_size), base_node(xc[1], k.dx_inv, k.domain_size),
                       base_node(xc[2], k.dx_inv, k.domain_size));
    T ctF[9];
    ctF[0] = F[0];
    ctF[1] = F[1];
    ctF[2] = (T(1) + dt * Cm[0]) * F[2] + dt * Cm[1] * F[5] + dt * Cm[2] * F[8];
    ctF[3] = F[3];
    ctF[4] = F[4];
    ctF[5] = dt * Cm[3] * F[2] + (T(1) + dt * Cm[4]) * F[5] + dt * Cm[5] * F[8];
    ctF[6] = F[6];
    ctF[7] = F[7];
    ctF[8] = dt * Cm[6] * F[2] + dt * Cm[7] * F[5] + (T(1) + dt * Cm[8]) * F[8];
    cloth_project_strain(k, ctF);

    // in-plane columns: ds * Dm^-1
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const T a = x1[d] - x0[d], b = x2[d] - x0[d];
      ctF[d * 3 + 0] = a * Dmi[0] + b * Dmi[2];
      ctF[d * 3 + 1] = a * Dmi[1] + b * Dmi[3];
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) Fd[(size_t)e * n_faces + j] = ctF[e];

    T VP[9];
    cloth_pk1(k, ctF, VP);
#pragma unroll
    for (int e = 0; e < 9; ++e) VP[e] *= volume;
    // tau = V * P[:,2] (x) F[:,2] is rank one: stored as its two factors, rows 0-2 and 3-5 of taus[6][N]
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      taus[(size_t)a * N + sf] = VP[a * 3 + 2];
      taus[(size_t)(3 + a) * N + sf] = ctF[a * 3 + 2];
    }
    // grad_N = Dm^-T * [-1 1 0; -1 0 1]
    T gN[6];
    gN[0] = Dmi[0] * T(-1) + Dmi[2] * T(-1);
    gN[1] = Dmi[0] * T(1) + Dmi[2] * T(0);
    gN[2] = Dmi[0] * T(0) + Dmi[2] * T(1);
    gN[3] = Dmi[1] * T(-1) + Dmi[3] * T(-1);
    gN[4] = Dmi[1] * T(1) + Dmi[3] * T(0);
    gN[5] = Dmi[1] * T(0) + Dmi[3] * T(1);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const T p0 = VP[d * 3 + 0], p1 = VP[d * 3 + 1];
      atomicAdd(&forces[(size_t)d * N + s0], -(p0 * gN[0] + p1 * gN[3]));
      atomicAdd(&forces[(size_t)d * N + s1], -(p0 * gN[1] + p1 * gN[4]));
      atomicAdd(&forces[(size_t)d * N + s2], -(p0 * gN[2] + p1 * gN[5]));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// P2G (replaces particle_to_grid_kernel :418-543 and clean_grid_kernel :585-602).
//
// One WARP per particle block (4^3 base cells).  The warp owns a private 6^3 x 4 shared-memory
// tile, so there are no atomics and no barriers, and the summation order is fixed (run-to-run
// reproducible).  Two alternating phases over batches of 32 particles of the block:
//   (a) lanes = particles: the raw particle record (prefetched into registers one batch ahead, so
//       the global loads are in flight during phase (b) of the previous batch) is turned into a
//       shared-memory row that makes the per-node work separable: for the 9 (i,j) stencil columns
//       W_ij = wx_i wy_j and u_ij,d = a'_d + Bdx_d0 i + Bdx_d1 j, plus Bdx_d2, mass and wz_k;
//   (b) lanes = the 27 stencil nodes (i,j,k): per particle 5 shared loads and 8 FMAs per lane:
//       wt = W_ij wz_k;  mass += wt m;  mom_d += wt (u_ij,d + Bdx_d2 k); accumulated in registers
//       over a run of particles with the same base cell (the order is sorted by cell key), then
//       added to the tile (distinct addresses across lanes -> plain read-modify-write).
// The tile is written to tiles[t][4][216]; grid_assemble_update sums the <= 8 tiles that overlap
// each grid block.  Touched grid blocks are appended to a device list (epoch-stamped, no flags scan).
// Work distribution: a device-side queue hands out blocks in two sweeps, blocks with more than the
// average population first (longest-processing-time-first keeps the tail short); the next work item
// and its first batch are fetched while the current block is still being processed.
// ------------------------------------------------------------------------------------------------
constexpr int kP2GThreads = 32;  // one warp per CTA: nothing is shared between warps

// Scalars per particle row: 9 x {A', G0, G1, G2} (per stencil column ij) + 3 x {z, zb0, zb1, zb2} (per k)
// = 48, padded to an odd multiple of 16 bytes (conflict-free 128-bit row stores, 16-byte aligned rows).
template <typename T> constexpr int kStageStride = 0;
template <> constexpr int kStageStride<double> = 50;  // 400 B = 25 x 16
template <> constexpr int kStageStride<float> = 52;   // 208 B = 13 x 16

template <typename T>
struct P2GSmem {
  T tile[4][kTileNodes];
  T stage[32][kStageStride<T>];
};

template <typename T>
struct alignas(16) Vec4 {
  T x, y, z, w;
};

// Raw record of one particle as P2G needs it (registers).
template <typename T>
struct P2GRaw {
  T p[3], v[3], vol, C[9], ex[6];
  bool face;  // carries a stress (taus): a cloth face (rank-one, two factors) or a volumetric particle (symmetric 3x3)
  bool solid; // volumetric particle (pid >= n_cloth): ex = V0 tau (xx, yy, zz, xy, xz, yz)
};

template <typename T>
__device__ __forceinline__ void p2g_load(P2GRaw<T>& r, bool valid, uint32_t s, uint32_t N, uint2 n_faces_cloth,
                                         const T* __restrict__ pos, const T* __restrict__ vel,
                                         const T* __restrict__ vol, const T* __restrict__ C,
                                         const int* __restrict__ pid, const T* __restrict__ forces,
                                         const T* __restrict__ taus) {
  if (!valid) return;
  const uint32_t p_id = (uint32_t)pid[s];
  r.solid = p_id >= n_faces_cloth.y;          // [faces | verts | volumetric particles]
  r.face = p_id < n_faces_cloth.x || r.solid;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    r.p[d] = pos[(size_t)d * N + s];
    r.v[d] = vel[(size_t)d * N + s];
  }
  r.vol = vol[s];
#pragma unroll
  for (int e = 0; e < 9; ++e) r.C[e] = C[(size_t)e * N + s];
  if (r.face) {  // stress lives on face particles only: tau = ex[0:3] (x) ex[3:6]
#pragma unroll
    for (int e = 0; e < 6; ++e) r.ex[e] = taus[(size_t)e * N + s];
  } else {       // nodal forces live on vertex particles only
#pragma unroll
    for (int d = 0; d < 3; ++d) r.ex[d] = forces[(size_t)d * N + s];
  }
}

// Phase (a) of P2G for one particle: node (i,j,k) receives
//   mass  += wx_i wy_j wz_k m                                   = A'_ij z_k
//   mom_d += wx_i wy_j wz_k (a'_d + Bx_d i + By_d j + Bz_d k)    = G_ij,d z_k + A'_ij zb_k,d
// with A'_ij = wx_i wy_j m, G_ij,d = wx_i wy_j (a'_d + Bx_d i + By_d j), z_k = wz_k,
// zb_k,d = wz_k k Bz_d / m  (B. = (m C - dt D^-1 tau) dx, a' = m v + m g dt + f dt - B. fx).
template <typename T>
__device__ __forceinline__ uint32_t p2g_derive(const P2GRaw<T>& r, T* row, T dt, const Consts<T>& k) {
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
  Vec4<T>* out = reinterpret_cast<Vec4<T>*>(row);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const T A = w[i][0] * w[j][1];
      Vec4<T> q;
      q.x = A * mass;
      q.y = A * (al[0] + Bx[0] * T(i) + By[0] * T(j));
      q.z = A * (al[1] + Bx[1] * T(i) + By[1] * T(j));
      q.w = A * (al[2] + Bx[2] * T(i) + By[2] * T(j));
      out[i * 3 + j] = q;
    }
#pragma unroll
  for (int kk = 0; kk < 3; ++kk) {
    const T zk = w[kk][2] * T(kk) * inv_mass;
    out[9 + kk] = Vec4<T>{w[kk][2], zk * Bz[0], zk * Bz[1], zk * Bz[2]};
  }
  // which neighbour blocks (+x,+y,+z combinations) does this stencil reach
  uint32_t reach = 1u;  // bit (dx*4+dy*2+dz) per reached neighbour; delta = 0 always
  if ((b[0] & 3u) >= 2u) reach |= reach << 4;
  if ((b[1] & 3u) >= 2u) reach |= reach << 2;
  if ((b[2] & 3u) >= 2u) reach |= reach << 1;
  // a massless particle (an inactive spare of the migration, csrc/migrate.cuh) deposits zeros: it must not make a
  // grid block "touched" -- the exchange plans the reachable blocks from the mass on the grid
  return mass > T(0) ? reach : 0u;
}

template <typename T>
__global__ void __launch_bounds__(kP2GThreads)
p2g_tile_kernel(uint32_t N, uint2 n_faces, const uint32_t* __restrict__ order,
                const uint32_t* __restrict__ sorted_key, const WorkRec* __restrict__ pb_work, uint32_t work_cap,
                const T* __restrict__ pos, const T* __restrict__ vel, const T* __restrict__ vol,
                const T* __restrict__ C, const int* __restrict__ pid, const T* __restrict__ forces,
                const T* __restrict__ taus, T* __restrict__ tiles, uint32_t* __restrict__ touched_epoch,
                uint32_t* __restrict__ touched_list, uint32_t epoch, DevCounters* ctr, T dt, Consts<T> k,
                uint32_t* __restrict__ nbr_list = nullptr) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  P2GSmem<T>& sm = *reinterpret_cast<P2GSmem<T>*>(smem_raw);
  const int lane = threadIdx.x;
  T* tile = &sm.tile[0][0];
  const uint32_t n_big = ctr->n_big, n_work = n_big + ctr->n_small;
  if (n_work == 0) return;
  const int ni = lane / 9, nj = (lane / 3) % 3, nk = lane % 3;  // node of this lane (lane < 27)
  const bool node_lane = lane < 27;
  const int ij = node_lane ? ni * 3 + nj : 0, kz = node_lane ? nk : 0;
  const uint32_t blocks_per_axis = (uint32_t)k.domain_size >> 2;

  // work item pipeline: `cur` is being processed, `nxt` is resolved, one queue ticket is in flight
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
  Item cur = resolve(ticket());
  if (cur.tile == kNone) return;
  Item nxt = resolve(ticket());
  // two-level prefetch: the slot index and key two batches ahead, the particle record one batch ahead
  // (a dependent load in program order would stall the warp for a full memory round trip)
  uint32_t s_a = 0, key_a = 0, s_b = 0, key_b = 0;  // _a: next batch to load records for; _b: the one after
  bool v_a = cur.begin + lane < cur.end;
  if (v_a) {
    s_a = order[cur.begin + lane];
    key_a = sorted_key[cur.begin + lane];
  }
  auto batch_after = [&](const Item& c, const Item& n, uint32_t p0, uint32_t& s, uint32_t& key) -> bool {
    // batch following the one that starts at p0 of item c (continuing into item n)
    uint32_t i;
    uint32_t end;
    if (p0 + 32 < c.end) {
      i = p0 + 32 + lane;
      end = c.end;
    } else if (n.tile != kNone) {
      i = n.begin + lane;
      end = n.end;
    } else {
      return false;
    }
    if (i >= end) return false;
    s = order[i];
    key = sorted_key[i];
    return true;
  };
  bool v_b = batch_after(cur, nxt, cur.begin, s_b, key_b);
  P2GRaw<T> raw;
  p2g_load(raw, v_a, s_a, N, n_faces, pos, vel, vol, C, pid, forces, taus);
  uint32_t key_cur = key_a;  // key of the particle whose record is in `raw`

  for (;;) {
    const uint32_t blk = cur.block;
    for (int i = lane; i < 4 * kTileNodes; i += 32) tile[i] = T(0);
    uint32_t reach = 0;
    T acc0 = T(0), acc1 = T(0), acc2 = T(0), acc3 = T(0);
    int run_cell = -1, last_cell = -1;
    uint32_t t_pending = ticket();  // the item after `nxt`
    Item nn{kNone, 0u, 0u, 0u};
    __syncwarp();

    for (uint32_t p0 = cur.begin; p0 < cur.end; p0 += 32) {
      const bool last_batch = p0 + 32 >= cur.end;
      // ---- phase (a): lanes = particles
      const bool valid = p0 + lane < cur.end;
      const int my_cell = (int)(key_cur & 63u);
      if (valid) reach |= p2g_derive(raw, &sm.stage[lane][0], dt, k);
      // run heads of this batch (the order is sorted by cell key inside a block)
      int prev = __shfl_up_sync(kFullMask, my_cell, 1);
      if (lane == 0) prev = last_cell;
      const uint32_t heads = __ballot_sync(kFullMask, valid && my_cell != prev);
      const int cnt = (int)min(32u, cur.end - p0);
      last_cell = __shfl_sync(kFullMask, my_cell, cnt - 1);
      __syncwarp();
      // ---- prefetch: records of the next batch, indices of the one after
      if (last_batch && nxt.tile != kNone) nn = resolve(t_pending);  // needed to look two batches ahead
      p2g_load(raw, v_b, s_b, N, n_faces, pos, vel, vol, C, pid, forces, taus);
      key_cur = key_b;
      {
        // the batch after next: relative to the batch that `raw` now holds
        const bool in_cur = !last_batch;
        const Item& c2 = in_cur ? cur : nxt;
        const Item& n2 = in_cur ? nxt : nn;
        const uint32_t p2 = in_cur ? p0 + 32 : nxt.begin;
        v_b = (in_cur || nxt.tile != kNone) ? batch_after(c2, n2, p2, s_b, key_b) : false;
      }
      // ---- phase (b): lanes = stencil nodes; the loads of particle q + 1 are issued before the math of q
      auto row_of = [&](int q) { return reinterpret_cast<const Vec4<T>*>(&sm.stage[q][0]); };
      Vec4<T> a4 = row_of(0)[ij], z4 = row_of(0)[9 + kz];
      for (int q = 0; q < cnt; ++q) {
        const int qn = q + 1 < cnt ? q + 1 : q;
        const Vec4<T> a4n = row_of(qn)[ij], z4n = row_of(qn)[9 + kz];
        if ((heads >> q) & 1u) {
          if (run_cell >= 0 && node_lane) {
            const int n = (((run_cell >> 4) & 3) + ni) * 36 + (((run_cell >> 2) & 3) + nj) * 6 + ((run_cell & 3) + nk);
            tile[0 * kTileNodes + n] += acc0;
            tile[1 * kTileNodes + n] += acc1;
            tile[2 * kTileNodes + n] += acc2;
            tile[3 * kTileNodes + n] += acc3;
          }
          __syncwarp();
          acc0 = acc1 = acc2 = acc3 = T(0);
          run_cell = __shfl_sync(kFullMask, my_cell, q);
        }
        acc0 += a4.x * z4.x;
        acc1 += a4.y * z4.x + a4.x * z4.y;
        acc2 += a4.z * z4.x + a4.x * z4.z;
        acc3 += a4.w * z4.x + a4.x * z4.w;
        a4 = a4n;
        z4 = z4n;
      }
      __syncwarp();
    }
    if (run_cell >= 0 && node_lane) {
      const int n = (((run_cell >> 4) & 3) + ni) * 36 + (((run_cell >> 2) & 3) + nj) * 6 + ((run_cell & 3) + nk);
      tile[0 * kTileNodes + n] += acc0;
      tile[1 * kTileNodes + n] += acc1;
      tile[2 * kTileNodes + n] += acc2;
      tile[3 * kTileNodes + n] += acc3;
    }
    __syncwarp();
    // ---- flush the tile (coalesced) and publish the touched grid blocks
    T* out = tiles + (size_t)cur.tile * (4 * kTileNodes);
    for (int i = lane; i < 4 * kTileNodes; i += 32) out[i] = tile[i];
    reach = __reduce_or_sync(kFullMask, reach);
    if (lane < 8 && ((reach >> lane) & 1u)) {
      uint32_t bx, by, bz;
      block_coords(blk, bx, by, bz);
      bx += (lane >> 2) & 1u;
      by += (lane >> 1) & 1u;
      bz += lane & 1u;
      if (bx < blocks_per_axis && by < blocks_per_axis && bz < blocks_per_axis) {
        const uint32_t g = morton3(bx, by, bz);
        if (atomicExch(&touched_epoch[g], epoch) != epoch) {
          const uint32_t idx = atomicAdd(&ctr->n_touched, 1u);
          touched_list[idx] = g;
          if (nbr_list) nbr_list[idx] = 0u;  // filled by the grid assembly (multi-GPU planning)
        }
      }
    }
    __syncwarp();
    if (nxt.tile == kNone) break;
    cur = nxt;
    nxt = nn;
  }
}

// ------------------------------------------------------------------------------------------------
// Grid: assemble the P2G tiles into the dense grid and/or update it
// (replaces gather_touched_grid_kernel :545-583 and update_grid_kernel :632-796).
// 64 threads (one per cell) x 4 grid blocks per CTA, grid-stride over the touched list.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void apply_boundary_and_colliders(const Consts<T>& k, int bc, uint32_t x, uint32_t y,
                                                             uint32_t z, T* v) {
  const uint32_t nb = (uint32_t)k.boundary_cells, dom = (uint32_t)k.domain_size;
  if (x < nb && v[0] < T(0)) v[0] = T(0);
  if (x >= dom - nb && v[0] > T(0)) v[0] = T(0);
  if (y < nb && v[1] < T(0)) v[1] = T(0);
  if (y >= dom - nb && v[1] > T(0)) v[1] = T(0);
  if (z < nb && v[2] < T(0)) v[2] = T(0);
  if (z >= dom - nb && v[2] > T(0)) v[2] = T(0);
  if (bc < 0) return;
  // hard-coded analytic colliders of the reference demos (:671-788)
  const T p[3] = {(T(x) + T(.5)) * k.dx, (T(y) + T(.5)) * k.dx, (T(z) + T(.5)) * k.dx};
  bool fixed = false, inside = false;
  T diff[3] = {T(0), T(0), T(0)}, n[3] = {T(0), T(0), T(0)}, dotnv = T(0);
  auto sphere = [&](T cx, T cy, T cz, T radius, bool need_approach) -> bool {
    n[0] = p[0] - cx; n[1] = p[1] - cy; n[2] = p[2] - cz;
    const T len = tsqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    const T dist = len - radius;
    const T inv = T(1) / len;
    n[0] *= inv; n[1] *= inv; n[2] *= inv;
    dotnv = T(0);
    if (dist < T(0)) {
      diff[0] = -v[0]; diff[1] = -v[1]; diff[2] = -v[2];   // collider velocity is zero
      dotnv = n[0] * diff[0] + n[1] * diff[1] + n[2] * diff[2];
      if (!need_approach || dotnv > T(0) || fixed) return true;
    }
    return false;
  };
  if (bc == 0) {
    inside = sphere(T(0.5), T(0.5), T(0.5), T(0.08), true);
  } else if (bc == 1) {
    fixed = true;
    inside = sphere(T(0.38), T(0.38), T(0.75), T(0.04), true);
    if (!inside) inside = sphere(T(0.38), T(0.62), T(0.75), T(0.04), true);  // only when outside the first (:711-733)
  } else if (bc == 2) {
    n[0] = T(0); n[1] = T(0); n[2] = T(1);
    if (p[2] - T(0.11) < T(0)) {
      inside = true;
      diff[0] = -v[0]; diff[1] = -v[1]; diff[2] = -v[2];
      dotnv = diff[2];
    }
  } else if (bc == 3) {
    fixed = true;
    const T span[2] = {T(0.3), T(0.7)};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (!inside) inside = sphere(span[q % 2], span[q / 2], T(0.5), T(0.02), false);
    }
  }
  if (inside) {
    if (fixed) {
      v[0] += diff[0]; v[1] += diff[1]; v[2] += diff[2];
    } else {
      const T frac = dotnv * (T(1) - k.sdf_friction);
#pragma unroll
      for (int d = 0; d < 3; ++d) v[d] += diff[d] * k.sdf_friction + n[d] * frac;
    }
  }
}

// HALO (multi-GPU, peer windows, halo.cuh): 1 = the assembly also SENDS -- it starts with the blocks of the shared-block
// list, stores them into their other sharers' windows and raises the delivery flags, then assembles the grid;
// 2 = the update also RECEIVES -- a listed block waits for its sharers' flags and adds their values in rank order
// before it is updated.  Saves the two exchange kernels and their launches between assembly and update.
template <typename T, bool ASSEMBLE, bool UPDATE, int HALO = 0>
__global__ void __launch_bounds__(kThreads)
grid_kernel(const uint32_t* __restrict__ touched_list, const DevCounters* __restrict__ ctr,
            const T* __restrict__ tiles, const uint2* __restrict__ pb_tile, const uint32_t* __restrict__ pb_epoch,
            uint32_t epoch, T* __restrict__ g_mass, T* __restrict__ g_vel, T* __restrict__ g_vstar, int bc,
            Consts<T> k, uint32_t* __restrict__ nbr_list = nullptr, const uint8_t* __restrict__ mark = nullptr,
            int want = -1, HaloFuse<T> hf = HaloFuse<T>{}) {
  // mark / want (multi-GPU): only blocks with mark[g] == want (0: not exchanged, 1: exchanged with other ranks)
  const uint32_t n_touched = ctr->n_touched;
  const uint32_t l = threadIdx.x & 63u;
  const uint32_t lx = (l >> 4) & 3u, ly = (l >> 2) & 3u, lz = l & 3u;
  // mass and momentum of this thread's cell of grid block (gx, gy, gz): sum of the P2G tiles that overlap it
  auto sum_tiles = [&](uint32_t gx, uint32_t gy, uint32_t gz, T& m, T* p) {
    m = T(0); p[0] = T(0); p[1] = T(0); p[2] = T(0);
    // resolve the <= 8 source particle blocks first (independent loads in flight together) ...
    uint2 tr[8];
    uint32_t node[8];
#pragma unroll
    for (int d = 0; d < 8; ++d) {
      const uint32_t ddx = (d >> 2) & 1u, ddy = (d >> 1) & 1u, ddz = d & 1u;
      // tile of particle block g - delta covers local cells [4 delta, 4 delta + 6): needs l < 2 where delta = 1
      const bool skip = (ddx && (lx >= 2u || gx == 0u)) || (ddy && (ly >= 2u || gy == 0u)) || (ddz && (lz >= 2u || gz == 0u));
      tr[d] = make_uint2(0u, 0u);
      node[d] = (lx + 4u * ddx) * 36u + (ly + 4u * ddy) * 6u + (lz + 4u * ddz);
      if (!skip) {
        const uint32_t src = morton3(gx - ddx, gy - ddy, gz - ddz);
        if (pb_epoch[src] == epoch) tr[d] = pb_tile[src];
      }
    }
    // ... then sum their chunks' tiles in a fixed order (delta, then chunk): reproducible
#pragma unroll
    for (int d = 0; d < 8; ++d) {
      const T* tl = tiles + (size_t)tr[d].x * (4 * kTileNodes) + node[d];
      for (uint32_t c = 0; c < tr[d].y; ++c, tl += 4 * kTileNodes) {
        m += tl[0];
        p[0] += tl[kTileNodes];
        p[1] += tl[2 * kTileNodes];
        p[2] += tl[3 * kTileNodes];
      }
    }
  };
  if (HALO == 1) {
    // The blocks of the shared-block list FIRST: summed, stored into their other sharers' windows, and the
    // delivery flags raised -- the system-scope fence and the NVLink round trip of that (~15 us) then run under
    // the assembly of the whole grid below, and the peers' data is there long before the update asks for it.
    // (The main loop sums these few blocks a second time, for the grid arrays and the neighbour masks.)
    bool sent = false;
    for (uint32_t i = blockIdx.x * 4u + (threadIdx.x >> 6); i < hf.n_list; i += gridDim.x * 4u) {
      const uint32_t g = hf.list[i];
      if (hf.touched_epoch[g] != hf.p2g_epoch) continue;  // not this rank's block now: nothing to send
      sent = true;
      uint32_t gx, gy, gz;
      block_coords(g, gx, gy, gz);
      T m, p[3];
      sum_tiles(gx, gy, gz, m, p);
      const uint32_t who = hf.sharers[i] & ~(1u << hf.rank);
      for (uint32_t q = 0; q < hf.world; ++q)
        if ((who >> q) & 1u) {
          T* d = halo_window_data<T>(hf.win.base[q], hf.xchg & 1u, hf.world, hf.rank, hf.cap, i);
          d[l] = m;
          d[64 + l] = p[0];
          d[128 + l] = p[1];
          d[192 + l] = p[2];
          if (l == 0) *halo_window_present(hf.win.base[q], hf.xchg & 1u, hf.world, hf.rank, hf.cap, i) = hf.xchg;
        }
    }
    halo_signal_peers_warp(hf.win, hf.rank, hf.world, hf.xchg, hf.peers, hf.done, sent);
  }
  for (uint32_t it = blockIdx.x * 4u + (threadIdx.x >> 6); it < n_touched; it += gridDim.x * 4u) {
    const uint32_t g = touched_list[it];
    if (want >= 0 && (int)mark[g] != want) continue;
    uint32_t gx, gy, gz;
    block_coords(g, gx, gy, gz);
    T m, p[3];
    const size_t cell = (size_t)g * 64 + l;
    const size_t vbase = (size_t)g * 192 + l;
    if (ASSEMBLE) {
      sum_tiles(gx, gy, gz, m, p);
      if (nbr_list) {
        // multi-GPU planning (halo.cuh): towards which of the 26 neighbour blocks does this block hold mass
        // in its outermost cell layer?  Only those can be reached by its particles in the next substep.
        uint32_t mk = 0;
        if (m > T(0)) {
          const uint32_t xb = 2u | (lx == 0u ? 1u : 0u) | (lx == 3u ? 4u : 0u);
          const uint32_t yb = 2u | (ly == 0u ? 1u : 0u) | (ly == 3u ? 4u : 0u);
          const uint32_t zb = 2u | (lz == 0u ? 1u : 0u) | (lz == 3u ? 4u : 0u);
#pragma unroll
          for (int d = 0; d < 27; ++d)
            if (((xb >> (d / 9)) & 1u) && ((yb >> ((d / 3) % 3)) & 1u) && ((zb >> (d % 3)) & 1u)) mk |= 1u << d;
        }
        mk = __reduce_or_sync(kFullMask, mk);
        if ((threadIdx.x & 31u) == 0u && mk) atomicOr(&nbr_list[it], mk);
      }
      if (!UPDATE) {
        g_mass[cell] = m;
        g_vel[vbase] = p[0];
        g_vel[vbase + 64] = p[1];
        g_vel[vbase + 128] = p[2];
      }
    } else {
      m = g_mass[cell];
      p[0] = g_vel[vbase];
      p[1] = g_vel[vbase + 64];
      p[2] = g_vel[vbase + 128];
      if (HALO == 2) {
        const uint32_t i = hf.index[g];
        if (i != kHaloNoIndex) {
          const uint32_t who = hf.sharers[i];
          unsigned char* window = hf.win.base[hf.rank];
          halo_wait_sharers(window, who, hf.rank, hf.world, hf.xchg, hf.err);  // both warps of the block's group
          T am = T(0), ap[3] = {T(0), T(0), T(0)};
          bool first = true;
          for (uint32_t q = 0; q < hf.world; ++q) {  // ascending rank order: the same bits on every sharer
            if (!((who >> q) & 1u)) continue;
            T vm = T(0), vp[3] = {T(0), T(0), T(0)};
            if (q == hf.rank) {
              vm = m; vp[0] = p[0]; vp[1] = p[1]; vp[2] = p[2];
            } else if (__ldcg(halo_window_present(window, hf.xchg & 1u, hf.world, q, hf.cap, i)) == hf.xchg) {
              const T* d = halo_window_data<T>(window, hf.xchg & 1u, hf.world, q, hf.cap, i);
              vm = __ldcg(d + l); vp[0] = __ldcg(d + 64 + l); vp[1] = __ldcg(d + 128 + l); vp[2] = __ldcg(d + 192 + l);
            }
            if (first) {
              am = vm; ap[0] = vp[0]; ap[1] = vp[1]; ap[2] = vp[2];
              first = false;
            } else {
              am += vm; ap[0] += vp[0]; ap[1] += vp[1]; ap[2] += vp[2];
            }
          }
          m = am; p[0] = ap[0]; p[1] = ap[1]; p[2] = ap[2];
          g_mass[cell] = m;  // the exchanged sums are what the grid holds from here on (getters, contact)
        }
      }
    }
    if (UPDATE) {
      if (m > T(0)) {
        p[0] /= m; p[1] /= m; p[2] /= m;
        apply_boundary_and_colliders(k, bc, (gx << 2) | lx, (gy << 2) | ly, (gz << 2) | lz, p);
        g_vstar[vbase] = p[0];
        g_vstar[vbase + 64] = p[1];
        g_vstar[vbase + 128] = p[2];
      }
      if (ASSEMBLE) g_mass[cell] = m;
      g_vel[vbase] = p[0];
      g_vel[vbase + 64] = p[1];
      g_vel[vbase + 128] = p[2];
    }
  }
}

// Zero the blocks that were touched by the previous P2G but not by this one (the reference zeroes
// all previously touched blocks in clean_grid_kernel; blocks touched again are overwritten here).
template <typename T>
__global__ void __launch_bounds__(kThreads)
clean_stale_blocks_kernel(const uint32_t* __restrict__ prev_list, const DevCounters* __restrict__ ctr,
                          const uint32_t* __restrict__ touched_epoch, uint32_t epoch, T* __restrict__ g_mass,
                          T* __restrict__ g_vel) {
  const uint32_t n_prev = ctr->n_touched_prev;
  const uint32_t l = threadIdx.x & 63u;
  for (uint32_t it = blockIdx.x * 4u + (threadIdx.x >> 6); it < n_prev; it += gridDim.x * 4u) {
    const uint32_t g = prev_list[it];
    if (touched_epoch[g] == epoch) continue;
    g_mass[(size_t)g * 64 + l] = T(0);
    g_vel[(size_t)g * 192 + l] = T(0);
    g_vel[(size_t)g * 192 + 64 + l] = T(0);
    g_vel[(size_t)g * 192 + 128 + l] = T(0);
  }
}

// ------------------------------------------------------------------------------------------------
// G2P + advection + reorder (replaces grid_to_particle_kernel<..,false> :798-924 and the
// floating-point part of compute_sorted_state_kernel :384-416).
//
// One WARP per particle block: the 6^3 x 3 velocity nodes the block's stencils can reach are staged
// in shared memory once, lanes = particles gather their 27 nodes from it.  Results are written to
// the OTHER particle buffer at the particle's position in the key-sorted order, so the physical
// order follows the spatial order with one step of lag and next step's loads stay coalesced.
// ------------------------------------------------------------------------------------------------
constexpr int kG2PWarps = 4;
// A/B switches (profiles/r02_g2p_ab.txt): velocity-tile buffers per warp (2 = the next chunk's tile is in flight while
// this one is gathered; 1 = half the shared memory, twice the resident CTAs) and the register cap via min CTAs per SM.
#ifndef GMPM_G2P_BUFS
#define GMPM_G2P_BUFS 2
#endif
#ifndef GMPM_G2P_MIN_CTAS
#define GMPM_G2P_MIN_CTAS 4
#endif
constexpr int kG2PBufs = GMPM_G2P_BUFS;

// 4- or 8-byte asynchronous global -> shared copy (LDGSTS); src_bytes = 0 writes zeros.
template <typename T>
__device__ __forceinline__ void cp_async_scalar(T* smem_dst, const T* gsrc, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? (int)sizeof(T) : 0;
  if (sizeof(T) == 8)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Software pipeline, two chunks deep: while chunk c is gathered, the velocity tile of chunk c+1 is in
// flight to the other shared-memory buffer (cp.async, no registers), its slot indices are loaded, and
// the queue ticket of chunk c+2 is outstanding; the first batch of positions of chunk c+1 is loaded
// during the last batch of chunk c.  No load is followed directly by its use.
template <typename T>
__global__ void __launch_bounds__(kG2PWarps * 32, GMPM_G2P_MIN_CTAS)
g2p_kernel(uint32_t N, const uint32_t* __restrict__ order, const WorkRec* __restrict__ pb_work, uint32_t work_cap,
           const T* __restrict__ g_vel, const T* __restrict__ pos, const T* __restrict__ vol,
           const int* __restrict__ pid, T* __restrict__ pos_out, T* __restrict__ vel_out, T* __restrict__ C_out,
           T* __restrict__ vol_out, int* __restrict__ pid_out, int* __restrict__ slot_of_pid,
           uint32_t* __restrict__ key_out, T* __restrict__ forces_out, DevCounters* ctr, T dt, Consts<T> k,
           const uint8_t* __restrict__ mark = nullptr, int want = -1, uint32_t n_cloth = 0xFFFFFFFFu) {
  // mark / want (multi-GPU): want = 0 -> only chunks whose stencils stay clear of the blocks exchanged with
  // other ranks (they can run while the exchange is in flight), want = 1 -> only the others, -1 -> all.
  __shared__ T s_vel[kG2PWarps][kG2PBufs][3][kTileNodes];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t n_big = ctr->n_big, n_work = n_big + ctr->n_small;
  const uint32_t blocks_per_axis = (uint32_t)k.domain_size >> 2;
  const T scale = T(4) * k.dx_inv;
  const T ca = (k.v_blend + T(1)) * T(.5), cb = (k.v_blend - T(1)) * T(.5);
  constexpr uint32_t kNone = 0xFFFFFFFFu;
  static_assert(kChunk <= 128 && kChunk > 96, "the slot-index rotation below assumes 4 batches per chunk");

  uint32_t* queue = want == 1 ? &ctr->work_grid : &ctr->work_g2p;  // the second part has a counter of its own
  auto ticket = [&]() -> uint32_t {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(queue, 1u);
    return t;
  };
  auto resolve = [&](uint32_t t_lane0) -> WorkRec {
    uint32_t t = __shfl_sync(kFullMask, t_lane0, 0);
    for (;;) {
      if (t >= n_work) return WorkRec{kNone, 0u, 0u, 0u};
      const WorkRec r = pb_work[t < n_big ? t : work_cap - 1u - (t - n_big)];
      if (want < 0) return r;
      // does the chunk reach an exchanged block?  lanes 0-7 look at the 8 blocks block + {0,1}^3
      uint32_t bx, by, bz;
      block_coords(r.block, bx, by, bz);
      bx += (lane >> 2) & 1u;
      by += (lane >> 1) & 1u;
      bz += lane & 1u;
      const bool hit = lane < 8 && bx < blocks_per_axis && by < blocks_per_axis && bz < blocks_per_axis &&
                       mark[morton3(bx, by, bz)] != 0;
      const int part = __any_sync(kFullMask, hit) ? 1 : 0;
      if (part == want) return r;
      t = __shfl_sync(kFullMask, ticket(), 0);  // not this launch's chunk: take the next one
    }
  };
  // velocity nodes the chunk's stencils can reach -> shared buffer `buf` (asynchronous)
  auto stage = [&](const WorkRec& w, int buf) {
    T* tvn = &s_vel[warp][buf][0][0];
    uint32_t bx, by, bz;
    block_coords(w.block, bx, by, bz);
    for (int n = lane; n < kTileNodes; n += 32) {
      const uint32_t tx = n / 36, ty = (n / 6) % 6, tz = n % 6;
      const uint32_t gx = bx + (tx >> 2), gy = by + (ty >> 2), gz = bz + (tz >> 2);
      const bool in = gx < blocks_per_axis && gy < blocks_per_axis && gz < blocks_per_axis;
      const size_t vb = in ? (size_t)morton3(gx, gy, gz) * 192 + (((tx & 3u) << 4) | ((ty & 3u) << 2) | (tz & 3u)) : 0;
      cp_async_scalar(tvn + n, g_vel + vb, in);
      cp_async_scalar(tvn + kTileNodes + n, g_vel + vb + 64, in);
      cp_async_scalar(tvn + 2 * kTileNodes + n, g_vel + vb + 128, in);
    }
  };
  auto load_sidx = [&](const WorkRec& w, uint32_t (&sidx)[4]) {
#pragma unroll
    for (int b4 = 0; b4 < 4; ++b4) {
      const uint32_t ii = w.begin + b4 * 32 + lane;
      sidx[b4] = ii < w.end ? order[ii] : 0u;
    }
  };

  // (tickets are taken one at a time here: resolve() may run through the queue to its end when it filters,
  // and a ticket taken earlier but not yet looked at would be lost)
  WorkRec w = resolve(ticket());
  if (w.tile == kNone) return;
  const uint32_t t1 = ticket();
  int buf = 0;
  uint32_t sidx[4], sidx_n[4] = {0u, 0u, 0u, 0u};
  stage(w, buf);
  cp_async_commit();
  load_sidx(w, sidx);
  WorkRec wn = resolve(t1);
  // first batch of positions of the first chunk
  T x[3] = {T(0), T(0), T(0)};
  if (w.begin + lane < w.end) {
    x[0] = pos[sidx[0]];
    x[1] = pos[(size_t)N + sidx[0]];
    x[2] = pos[(size_t)2 * N + sidx[0]];
  }
  for (;;) {
    uint32_t t2 = ticket();  // the chunk after next
    if (wn.tile != kNone) {
      if (kG2PBufs == 2) stage(wn, buf ^ 1);
      load_sidx(wn, sidx_n);
    }
    if (kG2PBufs == 2) {
      cp_async_commit();     // (possibly empty) group of the next chunk
      cp_async_wait<1>();    // this chunk's tile has landed
    } else {
      cp_async_wait<0>();    // single buffer: the tile was staged at the end of the previous chunk
    }
    __syncwarp();
    const T* tv = &s_vel[warp][buf][0][0];
    uint32_t i = w.begin + lane;
    uint32_t s = sidx[0];
    uint32_t s1 = sidx[1], s2 = sidx[2], s3 = sidx[3];
    const uint32_t n_batches = (w.end - w.begin + 31u) >> 5;  // warp-uniform trip count
    for (uint32_t bt = 0; bt < n_batches; ++bt) {
      const bool active = i < w.end;
      // prefetch the positions of the next batch: of this chunk, or the first one of the next chunk
      const bool last = bt + 1 == n_batches;
      const uint32_t i_n = last ? wn.begin + lane : i + 32;
      const uint32_t s_n = last ? sidx_n[0] : s1;
      const bool next_valid = last ? (wn.tile != kNone && i_n < wn.end) : (i_n < w.end);
      s1 = s2;
      s2 = s3;
      T x_n[3] = {T(0), T(0), T(0)};
      if (next_valid) {
        x_n[0] = pos[s_n];
        x_n[1] = pos[(size_t)N + s_n];
        x_n[2] = pos[(size_t)2 * N + s_n];
      }
      if (active) {
      const T pvol = vol[s];
      const int p = pid[s];
      uint32_t b[3];
      T fx[3], wt[3][3];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        b[d] = base_node(x[d], k.dx_inv, k.domain_size);
        fx[d] = x[d] * k.dx_inv - T(b[d]);
        wt[0][d] = T(0.5) * (T(1.5) - fx[d]) * (T(1.5) - fx[d]);
        wt[1][d] = T(0.75) - (fx[d] - T(1)) * (fx[d] - T(1));
        wt[2][d] = T(0.5) * (fx[d] - T(0.5)) * (fx[d] - T(0.5));
      }
      const int n0 = (int)((b[0] & 3u) * 36u + (b[1] & 3u) * 6u + (b[2] & 3u));
      // Separable accumulation of nv[a] = sum w v_a and the first moments M[a][c] = sum w v_a n_c
      // (n = stencil offset 0..2 per axis); then sum w v_a (n_c - fx_c) = M[a][c] - fx_c nv[a].
      const T wz2x2 = T(2) * wt[2][2];
      T nv[3] = {T(0), T(0), T(0)}, Mx[3] = {T(0), T(0), T(0)}, My[3] = {T(0), T(0), T(0)}, Mz[3] = {T(0), T(0), T(0)};
#pragma unroll
      for (int ii = 0; ii < 3; ++ii) {
        T s00[3] = {T(0), T(0), T(0)}, s0y[3] = {T(0), T(0), T(0)}, s0z[3] = {T(0), T(0), T(0)};
#pragma unroll
        for (int jj = 0; jj < 3; ++jj) {
          const int n = n0 + ii * 36 + jj * 6;
          const T wy = wt[jj][1];
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            const T v0 = tv[a * kTileNodes + n], v1 = tv[a * kTileNodes + n + 1], v2 = tv[a * kTileNodes + n + 2];
            const T t0 = wt[0][2] * v0 + wt[1][2] * v1 + wt[2][2] * v2;
            const T t1 = wt[1][2] * v1 + wz2x2 * v2;
            s00[a] += wy * t0;
            s0z[a] += wy * t1;
            if (jj) s0y[a] += (T(jj) * wy) * t0;
          }
        }
        const T wx = wt[ii][0];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          nv[a] += wx * s00[a];
          My[a] += wx * s0y[a];
          Mz[a] += wx * s0z[a];
          if (ii) Mx[a] += (T(ii) * wx) * s00[a];
        }
      }
      // an inactive spare of the migration (volumetric, rest volume 0) does not follow the flow: it keeps its
      // parking point, with zero velocity and affine matrix, so the spares stay scattered and inside the blocks
      // this rank exchanges
      if ((uint32_t)p >= n_cloth && !(pvol > T(0))) {
#pragma unroll
        for (int a = 0; a < 3; ++a) nv[a] = Mx[a] = My[a] = Mz[a] = T(0);
      }
      T S[9];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        S[a * 3 + 0] = Mx[a] - fx[0] * nv[a];
        S[a * 3 + 1] = My[a] - fx[1] * nv[a];
        S[a * 3 + 2] = Mz[a] - fx[2] * nv[a];
      }
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int c = 0; c < 3; ++c)
          // cloth: the reference's blended affine matrix (:905-915); volumetric particles keep grad v itself
          C_out[(size_t)(a * 3 + c) * N + i] = (uint32_t)p >= n_cloth ? scale * S[a * 3 + c]
                                                                      : ca * (scale * S[a * 3 + c]) + cb * (scale * S[c * 3 + a]);
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        x[d] += nv[d] * dt;
        vel_out[(size_t)d * N + i] = nv[d];
        pos_out[(size_t)d * N + i] = x[d];
        forces_out[(size_t)d * N + i] = T(0);  // the next constitutive update accumulates into zeros
      }
      vol_out[i] = pvol;
      pid_out[i] = p;
      slot_of_pid[p] = (int)i;
      key_out[i] = cell_key(base_node(x[0], k.dx_inv, k.domain_size), base_node(x[1], k.dx_inv, k.domain_size),
                            base_node(x[2], k.dx_inv, k.domain_size));
      }  // active
      i = i_n;
      s = s_n;
      x[0] = x_n[0];
      x[1] = x_n[1];
      x[2] = x_n[2];
    }
    __syncwarp();  // every lane is done with this tile before the buffer is staged again
    if (wn.tile == kNone) break;
    if (kG2PBufs == 1) {     // single buffer: stage the next chunk's tile now that this one has been consumed
      stage(wn, 0);
      cp_async_commit();
    }
    w = wn;
#pragma unroll
    for (int b4 = 0; b4 < 4; ++b4) sidx[b4] = sidx_n[b4];
    wn = resolve(t2);
    if (kG2PBufs == 2) buf ^= 1;
  }
  cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// Layout converters for the host-facing getters (reference layout, visible particle order).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void gather_particle_field_kernel(uint32_t N, int comps, const T* __restrict__ src,
                                             const int* __restrict__ pid_of_vis, const int* __restrict__ slot_of_pid,
                                             T* __restrict__ out) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const int s = slot_of_pid[pid_of_vis[v]];
  for (int c = 0; c < comps; ++c) out[(size_t)v * comps + c] = src[(size_t)c * N + s];
}
// Reference layout of taus (9 per particle, zero on vertex particles) from the two stored factors.
template <typename T>
__global__ void gather_taus_kernel(uint32_t N, uint32_t n_faces, const T* __restrict__ taus,
                                   const int* __restrict__ pid_of_vis, const int* __restrict__ slot_of_pid,
                                   T* __restrict__ out, uint32_t n_cloth = 0xFFFFFFFFu) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const int p = pid_of_vis[v];
  const int s = slot_of_pid[p];
  if ((uint32_t)p >= n_cloth) {  // volumetric particle: symmetric V0 tau stored as (xx, yy, zz, xy, xz, yz)
    const T t[6] = {taus[s], taus[(size_t)N + s], taus[(size_t)2 * N + s], taus[(size_t)3 * N + s],
                    taus[(size_t)4 * N + s], taus[(size_t)5 * N + s]};
    T* o = out + (size_t)v * 9;
    o[0] = t[0]; o[4] = t[1]; o[8] = t[2];
    o[1] = o[3] = t[3];
    o[2] = o[6] = t[4];
    o[5] = o[7] = t[5];
    return;
  }
  const bool face = (uint32_t)p < n_faces;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b)
      out[(size_t)v * 9 + a * 3 + b] = face ? taus[(size_t)a * N + s] * taus[(size_t)(3 + b) * N + s] : T(0);
}
// Face getters: record j of the (spatially ordered) face arrays holds face id fid[j].
template <typename T>
__global__ void transpose_planar_kernel(uint32_t n, int comps, const T* __restrict__ src,
                                        const int* __restrict__ fid, T* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t o = fid ? (size_t)fid[i] : (size_t)i;
  for (int c = 0; c < comps; ++c) out[o * comps + c] = src[(size_t)c * n + i];
}

// Face-record refresh: sort key of record j = current slot of its face particle ...
__global__ void face_slot_keys_kernel(uint32_t n_faces, const int* __restrict__ fid,
                                      const int* __restrict__ slot_of_pid, uint32_t* __restrict__ keys) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n_faces) keys[j] = (uint32_t)slot_of_pid[fid[j]];
}
// ... and the gather of the records into that order (perm[j] = old record index).
template <typename T>
__global__ void permute_faces_kernel(uint32_t n_faces, const uint32_t* __restrict__ perm,
                                     const int* __restrict__ fid_in, const int* __restrict__ tri_in,
                                     const T* __restrict__ F_in, const T* __restrict__ Dm_in, int* __restrict__ fid_out,
                                     int* __restrict__ tri_out, T* __restrict__ F_out, T* __restrict__ Dm_out) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_faces) return;
  const uint32_t s = perm[j];
  fid_out[j] = fid_in[s];
#pragma unroll
  for (int c = 0; c < 3; ++c) tri_out[(size_t)c * n_faces + j] = tri_in[(size_t)c * n_faces + s];
#pragma unroll
  for (int e = 0; e < 9; ++e) F_out[(size_t)e * n_faces + j] = F_in[(size_t)e * n_faces + s];
#pragma unroll
  for (int e = 0; e < 4; ++e) Dm_out[(size_t)e * n_faces + j] = Dm_in[(size_t)e * n_faces + s];
}
template <typename T>
__global__ void grid_vec_to_reference_kernel(uint32_t n_cells, const T* __restrict__ src, T* __restrict__ out) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cells) return;
  const size_t vb = (size_t)(c >> 6) * 192 + (c & 63u);
  out[(size_t)c * 3 + 0] = src[vb];
  out[(size_t)c * 3 + 1] = src[vb + 64];
  out[(size_t)c * 3 + 2] = src[vb + 128];
}
__global__ void flags_from_list_kernel(const uint32_t* __restrict__ list, const DevCounters* __restrict__ ctr,
                                       uint32_t* __restrict__ flags) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ctr->n_touched) flags[list[i]] = 1u;
}
// DumpCpuState (cuda_mpm_model.cu:245-265): vertex positions in original vertex order.
template <typename T>
__global__ void gather_vertex_positions_kernel(uint32_t n_verts, uint32_t n_faces, uint32_t N,
                                               const T* __restrict__ pos, const int* __restrict__ slot_of_pid,
                                               T* __restrict__ out) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_verts) return;
  const int s = slot_of_pid[n_faces + v];
  out[(size_t)v * 3 + 0] = pos[s];
  out[(size_t)v * 3 + 1] = pos[N + s];
  out[(size_t)v * 3 + 2] = pos[2 * N + s];
}

}  // namespace gmpm
