// halo.cuh -- shared-block exchange for scenes partitioned across GPUs (one process / state per GPU).
//
// The reference is single-GPU (settings.h:40 G_DEVICE_COUNT = 1, never used); this is the multi-GPU
// extension SURVEY.md 8(e) plans.  Cloth particles are owned by MESH (a triangle gathers its three
// vertices, so a mesh never straddles ranks); every rank addresses the same block-major grid.  The
// only coupling between ranks is through grid blocks that particles of more than one rank reach, so
// the exchange step sits between ParticleToGrid and UpdateGrid:
//
//   plan (for the NEXT substep, off the critical path, on a side stream):
//   1. halo_flags   : flags[b] = 1 if block b was touched by this rank's P2G NOW, or is the neighbour of a
//                     touched block that holds mass in its cell layer facing b.  A particle moves less
//                     than one CELL per substep (CFL), so next substep's stencils stay inside the cells
//                     with mass now dilated by one cell: the flagged set contains every block the rank
//                     can touch in the next substep.
//      -- host: all-reduce SUM of flags over the ranks (NCCL) --
//   2. halo_plan    : shared blocks = {b : flags[b] > 1}, listed in ascending block id by a
//                     deterministic two-kernel compaction (every rank derives the SAME list from the
//                     same reduced flags); the count goes to a pinned host word, asynchronously
//   exchange (this substep, between ParticleToGrid and UpdateGrid, with the list planned one substep ago):
//   3. halo_pack    : packed[i][4][64] = this rank's mass and momentum of shared block i (zeros where
//                     its P2G did not touch it); the host already knows the count -> no read-back
//      -- host: all-reduce SUM of packed[0 : n_shared * 256] --
//   4. halo_unpack  : the summed mass / momentum are written back into the blocks this rank touched.
//   The containment assumption is CHECKED on the device: a touched block outside the previous dilated
//   set bumps an error counter that the next halo_pack turns into an error.
//
// UpdateGrid then runs on each rank's own touched list: shared blocks are updated redundantly from
// identical sums, so all owners see bit-identical grid velocities, and G2P stays local.  No particle
// changes owner (ownership is by mesh), so there is no migration step for cloth scenes.
#pragma once
#include "kernels.cuh"

namespace gmpm {

constexpr int kHaloThreads = 256;
constexpr int kHaloBlocksPerCta = 1024;  // grid blocks scanned by one CTA (4 per thread)

// One thread per (touched block, neighbour): flags[neighbour] = 1 (flags were zeroed).  The centre thread
// also checks the block against the previous dilated set `own_prev` (containment / CFL check).
__global__ void halo_flags_kernel(const uint32_t* __restrict__ touched_list, const DevCounters* __restrict__ ctr,
                                  uint32_t blocks_per_axis, const uint8_t* __restrict__ own_prev,
                                  const uint32_t* __restrict__ nbr_list, uint8_t* __restrict__ flags,
                                  uint32_t* __restrict__ err, uint8_t contribution) {
  const uint32_t n = ctr->n_touched * 27u;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t g = touched_list[i / 27u], d = i % 27u;
    // nbr_list (from the grid assembly): neighbour directions in which the block has mass in its outer
    // cell layer; without it every neighbour is flagged
    if (nbr_list && d != 13u && !((nbr_list[i / 27u] >> d) & 1u)) continue;
    uint32_t bx, by, bz;
    block_coords(g, bx, by, bz);
    const uint32_t nx = bx + d / 9u - 1u, ny = by + (d / 3u) % 3u - 1u, nz = bz + d % 3u - 1u;  // wraps below 0
    if (nx < blocks_per_axis && ny < blocks_per_axis && nz < blocks_per_axis) flags[morton3(nx, ny, nz)] = contribution;
    if (d == 13u && own_prev && own_prev[g] == 0) atomicAdd(err, 1u);
  }
}

// Shared = reached by more than one rank.  The ranks contribute either 1 each (flags = number of sharers) or,
// in peer mode (<= 8 ranks), the bit 1 << rank (flags = sharer set; the SUM all-reduce of distinct bits is their OR).
__device__ __forceinline__ bool halo_is_shared(uint32_t f, bool bits) { return bits ? (f & (f - 1u)) != 0u : f > 1u; }
__device__ __forceinline__ uint32_t halo_shared_mask(const uint8_t* __restrict__ flags, uint32_t n_blocks,
                                                     uint32_t b0, bool bits) {
  uint32_t m = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (b0 + j < n_blocks && halo_is_shared(flags[b0 + j], bits)) m |= 1u << j;
  return m;
}

// Pass 1: number of shared blocks in each CTA's span.
__global__ void __launch_bounds__(kHaloThreads)
halo_count_kernel(uint32_t n_blocks, const uint8_t* __restrict__ flags, uint32_t* __restrict__ cta_counts, bool bits,
                  uint32_t* __restrict__ union_out) {
  __shared__ uint32_t s_sum;
  if (threadIdx.x == 0) s_sum = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) *union_out = 0;  // accumulated by the compaction that follows
  __syncthreads();
  const uint32_t b0 = (blockIdx.x * kHaloThreads + threadIdx.x) * 4u;
  uint32_t c = __popc(halo_shared_mask(flags, n_blocks, b0, bits));
  c = __reduce_add_sync(kFullMask, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_sum, c);
  __syncthreads();
  if (threadIdx.x == 0) cta_counts[blockIdx.x] = s_sum;
}

// Pass 2: ascending list of shared blocks (stable: CTA offset = sum of the preceding CTA counts,
// in-CTA offset by an exclusive scan over the threads).
__global__ void __launch_bounds__(kHaloThreads)
halo_compact_kernel(uint32_t n_blocks, const uint8_t* __restrict__ flags, const uint32_t* __restrict__ cta_counts,
                    uint32_t* __restrict__ list, uint32_t list_cap, uint32_t* __restrict__ n_shared, bool bits,
                    uint8_t* __restrict__ sharers, uint32_t* __restrict__ union_out, uint32_t rank) {
  __shared__ uint32_t s_warp[kHaloThreads / 32];
  __shared__ uint32_t s_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // offset of this CTA
  uint32_t pre = 0;
  for (uint32_t i = threadIdx.x; i < blockIdx.x; i += kHaloThreads) pre += cta_counts[i];
  pre = __reduce_add_sync(kFullMask, pre);
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  if (lane == 0 && pre) atomicAdd(&s_base, pre);
  const uint32_t b0 = (blockIdx.x * kHaloThreads + threadIdx.x) * 4u;
  const uint32_t m = halo_shared_mask(flags, n_blocks, b0, bits);
  const uint32_t c = __popc(m);
  uint32_t incl = c;
  uint32_t who = 0;  // ranks this rank shares one of this thread's blocks with (peer mode)
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
  for (int j = 0; j < 4; ++j)
    if ((m >> j) & 1u) {
      if (off < list_cap) {
        list[off] = b0 + j;
        if (bits) {
          sharers[off] = flags[b0 + j];
          if ((flags[b0 + j] >> rank) & 1u) who |= flags[b0 + j];  // q in peers(r) <=> r in peers(q)
        }
      }
      ++off;
    }
  if (bits) {
    who = __reduce_or_sync(kFullMask, who);
    if (lane == 0 && who) atomicOr(union_out, who);
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kHaloThreads - 1) *n_shared = off;  // grand total
}

// packed[i][c][l], c = 0: mass, 1..3: momentum.  One CTA-quarter (64 threads) per (block, channel).
template <typename T, bool UNPACK>
__global__ void __launch_bounds__(kHaloThreads)
halo_pack_kernel(uint32_t n, const uint32_t* __restrict__ list, const uint32_t* __restrict__ touched_epoch,
                 uint32_t epoch, T* __restrict__ g_mass, T* __restrict__ g_vel, T* __restrict__ packed) {
  const uint32_t c = threadIdx.x >> 6, l = threadIdx.x & 63u;
  for (uint32_t i = blockIdx.x; i < n; i += gridDim.x) {
    const uint32_t g = list[i];
    const bool own = touched_epoch[g] == epoch;
    T* cell = c == 0 ? g_mass + (size_t)g * 64 + l : g_vel + (size_t)g * 192 + (size_t)(c - 1) * 64 + l;
    T* slot = packed + (size_t)i * 256 + threadIdx.x;
    if (UNPACK) {
      if (own) *cell = *slot;
    } else {
      *slot = own ? *cell : T(0);
    }
  }
}

// index[list[i]] = i (set) or kHaloNoIndex (clear) for the first min(*n, cap) entries of a planned list: the
// grid block -> list position map the fused exchange (grid_kernel HALO = 1 / 2) looks its blocks up in.
__global__ void halo_index_kernel(const uint32_t* __restrict__ n_ptr, uint32_t cap, const uint32_t* __restrict__ list,
                                  uint32_t* __restrict__ index, bool set) {
  const uint32_t n = min(*n_ptr, cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    index[list[i]] = set ? i : kHaloNoIndex;
}

// mark[list[i]] = value for the first n entries (the blocks of the current exchange).
__global__ void halo_mark_kernel(uint32_t n, const uint32_t* __restrict__ list, uint8_t value,
                                 uint8_t* __restrict__ mark) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) mark[list[i]] = value;
}

// ------------------------------------------------------------------------------------------------
// Peer mode: the exchange itself over NVLink peer memory, no collective on the critical path.
//
// Every rank owns a receive WINDOW (one cudaMalloc, opened by the other ranks through CUDA IPC, or shared
// by pointer when several states live in one process):
//     uint32 flag[64]                      flag[r] = number of the last exchange rank r has delivered
//     uint32 present[2][world][cap]        present[e & 1][r][i] = e if rank r touched block i in exchange e
//     T      data[2][world][cap][4][64]    data[e & 1][r][i] = rank r's mass / momentum of shared block i
// i indexes the shared-block list, which is identical on all ranks, so no per-pair lists are needed.
//   send : for every listed block, this rank's values (zeros if it did not touch the block) are STORED into the
//          window of every other sharer; after a system-scope fence the last CTA stores the exchange number
//          into flag[rank] of every rank it shares a block with (release).
//   recv : spins (acquire) until every rank it shares a block with has delivered this exchange, then sums the
//          contributions of a block's sharers IN ASCENDING RANK ORDER -- every sharer computes the same
//          bits -- into the blocks this rank touched.
// The two window halves alternate: a rank can be at most one exchange ahead of a neighbour (its next recv
// needs the neighbour's next send, which follows the neighbour's current recv in stream order).
// ------------------------------------------------------------------------------------------------
// One CTA per listed block.  A block this rank did not touch sends nothing: its `present` word keeps an older
// exchange number and the receivers read it as zeros.
template <typename T>
__global__ void __launch_bounds__(kHaloPeerThreads)
halo_peer_send_kernel(uint32_t n, const uint32_t* __restrict__ list, const uint8_t* __restrict__ sharers,
                      const uint32_t* __restrict__ touched_epoch, uint32_t epoch, T* __restrict__ g_mass,
                      T* __restrict__ g_vel, HaloWindows win, uint32_t rank, uint32_t world, uint32_t cap,
                      uint32_t xchg, const uint32_t* __restrict__ union_mask, uint32_t* __restrict__ done) {
  using P2 = typename Pair<T>::type;
  const uint32_t i = blockIdx.x;
  const uint32_t g = list[i];
  if (touched_epoch[g] == epoch) {
    const uint32_t who = sharers[i] & ~(1u << rank);
    const P2 v = *halo_block_pair<T>(g_mass, g_vel, g, threadIdx.x);
    for (uint32_t q = 0; q < world; ++q)
      if ((who >> q) & 1u) {
        reinterpret_cast<P2*>(halo_window_data<T>(win.base[q], xchg & 1u, world, rank, cap, i))[threadIdx.x] = v;
        if (threadIdx.x == 0) *halo_window_present(win.base[q], xchg & 1u, world, rank, cap, i) = xchg;
      }
  }
  halo_signal_peers(win, rank, world, xchg, union_mask, done);
}

template <typename T>
__global__ void __launch_bounds__(kHaloPeerThreads)
halo_peer_recv_kernel(uint32_t n, const uint32_t* __restrict__ list, const uint8_t* __restrict__ sharers,
                      const uint32_t* __restrict__ touched_epoch, uint32_t epoch, T* __restrict__ g_mass,
                      T* __restrict__ g_vel, unsigned char* __restrict__ window, uint32_t rank, uint32_t world,
                      uint32_t cap, uint32_t xchg, uint32_t* __restrict__ err) {
  using P2 = typename Pair<T>::type;
  const uint32_t i = blockIdx.x;
  const uint32_t g = list[i];
  if (touched_epoch[g] != epoch) return;  // not this rank's block now (the list is the dilated set)
  const uint32_t who = sharers[i];
  // wait for the deliveries of this block's other sharers
  if (threadIdx.x < 32) halo_wait_sharers(window, who, rank, world, xchg, err);
  __syncthreads();
  P2* cell = halo_block_pair<T>(g_mass, g_vel, g, threadIdx.x);
  P2 acc;
  bool first = true;
  for (uint32_t q = 0; q < world; ++q) {  // ascending rank order: the same bits on every sharer
    if (!((who >> q) & 1u)) continue;
    P2 v;
    if (q == rank) {
      v = *cell;
    } else if (__ldcg(halo_window_present(window, xchg & 1u, world, q, cap, i)) == xchg) {
      v = __ldcg(reinterpret_cast<const P2*>(halo_window_data<T>(window, xchg & 1u, world, q, cap, i)) + threadIdx.x);
    } else {
      v.x = v.y = T(0);
    }
    if (first) {
      acc = v;
      first = false;
    } else {
      acc.x += v.x;
      acc.y += v.y;
    }
  }
  *cell = acc;
}

}  // namespace gmpm
