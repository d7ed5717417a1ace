// sort.cuh -- stable LSD radix sort of (key, value) uint32 pairs on key bits [0, num_bit), and the
// particle binning built on it, as ONE cooperative persistent kernel.
//
// Replaces the reference's radix_sort (radix_sort.cuh:202-286: 2-bit digits, Hillis-Steele block
// scans, recursive Blelloch scan: 8 digit passes x ~5 global sweeps x 3+ launches for its 16 key bits).
// Here the digits are up to 11 bits wide, so the reference's 16 visible key bits are 2 passes and a
// full 21-bit cell key is 2 passes (27 bits at 512^3: 3), and all passes run inside one launch with
// grid-wide barriers between their phases:
//   hist    : per-tile digit histogram (warp-aggregated shared-memory counting) -> hist[tile][digit]
//   scan    : one warp per digit: exclusive scan over the tiles, digit totals
//   scatter : stable ranking inside the tile with warp match-any, scatter to the other buffer
// and, when the sort is the particle binning, two more phases (block ranges + work records, see
// kernels.cuh).  Stability (ties keep input order) is what makes the result bit-identical to the
// reference's permutation and the summation order of P2G reproducible.
// Traffic per pass: 4 B (hist) + 8 B read + 8 B write per element; at the 1M-particle scene everything
// is L2 resident and the cost is the 3 barriers per pass.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace gmpm {

constexpr int kMaxRadixBits = 11;
constexpr int kMaxRadix = 1 << kMaxRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
#ifndef GMPM_SORT_ITEMS
#define GMPM_SORT_ITEMS 16
#endif
constexpr int kSortItemsPerLane = GMPM_SORT_ITEMS;                                  // per pass-through of a warp
constexpr int kSortWarpChunk = 32 * kSortItemsPerLane;                 // 512 consecutive items / warp
constexpr int kSortTile = kSortWarps * kSortWarpChunk;                 // 4096 items / CTA
// Small inputs use HALF the items per lane (tile = 2048): with few tiles the phases are pure latency chains and a
// shorter chain per tile wins (measured: binning of the 29 602-particle smoke scene 92 -> 77 us; at 1M particles the
// full tile is better, 88 against 98 us).  The loops stay unrolled over kSortItemsPerLane and leave early.
constexpr int kSortItemsSmall = kSortItemsPerLane / 2;
constexpr uint32_t kSortSmallN = 131072;
__host__ __device__ inline uint32_t sort_items_for(uint32_t n) { return n < kSortSmallN ? kSortItemsSmall : kSortItemsPerLane; }
constexpr size_t kSortSmemBytes = sizeof(uint32_t) * (size_t)(kSortWarps + 1) * kMaxRadix + 64;

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

struct SortPass {
  int shift, bits;
};

struct SortArgs {
  const uint32_t* keys_in;   // input keys (pass 0 reads them; never written)
  const uint32_t* vals_in;   // input values, or nullptr: values of pass 0 are 0..n-1
  uint32_t* keys[2];         // ping-pong buffers: pass p writes keys[p & 1]
  uint32_t* vals[2];
  uint32_t n, n_tiles;
  uint32_t items;            // items per lane and pass-through of a warp (sort_items_for(n)); tile = 8 warps x 32 x items
  int n_passes;
  SortPass pass[4];
  uint32_t* hist;            // [radix][n_tiles] scratch (digit-major: the scan reads contiguous rows)
  uint32_t* totals;          // [radix] scratch
};

inline size_t radix_sort_scratch_words(size_t n) {
  // sized for the SMALL tile whatever n is: the same scratch also serves sorts of fewer items (the face records)
  const size_t tile = (size_t)kSortWarps * 32 * kSortItemsSmall;
  const size_t n_tiles = (n + tile - 1) / tile;
  return (size_t)kMaxRadix * (n_tiles ? n_tiles : 1) + kMaxRadix;
}

// Splits num_bit key bits into the fewest passes of <= kMaxRadixBits, evenly.
__host__ __device__ inline int plan_sort_passes(int num_bit, SortPass* pass) {
  if (num_bit <= 0) return 0;
  const int n = (num_bit + kMaxRadixBits - 1) / kMaxRadixBits;
  int shift = 0;
  for (int p = 0; p < n; ++p) {
    const int bits = (num_bit - shift + (n - p) - 1) / (n - p);
    pass[p] = SortPass{shift, bits};
    shift += bits;
  }
  return n;
}

// Optional key window (binning only): the keys are cell keys; sorting them RELATIVE to the block-aligned
// corner (ox, oy, oz) of the particles' bounding box needs 3 w + 6 bits instead of 3 domain_bits, which
// saves a whole pass on the 256^3 and 512^3 grids of the multi-GPU scenes.  The relative key keeps the
// block grouping and the cell order inside a block, which is all the binning needs.
struct SortWindow {
  bool on;
  uint32_t ox, oy, oz;   // window origin in block coordinates, DILATED onto the Morton bit positions of its axis
  int n_passes;          // pass plan for the relative keys (decided on the device)
  SortPass pass[4];
  int final_buf;         // buffer pair the result must end up in (the host's expectation)
  uint32_t* stamp_out;   // instrumentation: ns since t0 at the end of hist / scan / scatter of pass 0 (or null)
  unsigned long long t0;
  // A window PREDICTED from the previous sort (its bounding box grown by one block) is not guaranteed to hold
  // every key: pass 0 checks each relative key against rel_bits and raises *violation; after the barrier that
  // follows its histogram the whole grid then falls back to the absolute keys (on = false on return).
  int rel_bits;
  uint32_t* violation;   // null for an exact window
};
// Morton bit positions of the three axes inside a block id (x most significant of each triple).
constexpr uint32_t kMortonZ = 0x09249249u, kMortonY = kMortonZ << 1, kMortonX = kMortonZ << 2;
// Per-axis subtraction / addition directly on the interleaved fields (dilated-integer arithmetic).
__device__ __forceinline__ uint32_t morton_sub(uint32_t b, uint32_t ox, uint32_t oy, uint32_t oz) {
  // borrows run through the zero gaps of the dilated operands
  return (((b & kMortonX) - ox) & kMortonX) | (((b & kMortonY) - oy) & kMortonY) | (((b & kMortonZ) - oz) & kMortonZ);
}
__device__ __forceinline__ uint32_t morton_add(uint32_t b, uint32_t ox, uint32_t oy, uint32_t oz) {
  return ((((b & kMortonX) | ~kMortonX) + ox) & kMortonX) | ((((b & kMortonY) | ~kMortonY) + oy) & kMortonY) |
         ((((b & kMortonZ) | ~kMortonZ) + oz) & kMortonZ);
}
__device__ __forceinline__ uint32_t window_key(uint32_t key, uint32_t ox, uint32_t oy, uint32_t oz) {
  return (morton_sub(key >> 6, ox, oy, oz) << 6) | (key & 63u);
}
__device__ __forceinline__ uint32_t window_block_abs(uint32_t rel_block, const SortWindow& w) {
  return morton_add(rel_block, w.ox, w.oy, w.oz);
}

// One full sort; every thread of the cooperative grid calls it.  `smem` = kSortSmemBytes.
// Returns the index (0 or 1) of the buffer pair that holds the result.
__device__ __forceinline__ int radix_sort_device(const SortArgs& a, uint32_t* smem,
                                                 cooperative_groups::grid_group& grid, SortWindow* win = nullptr) {
  const uint32_t items = a.items;  // items per lane (tile size): uniform over the grid
  bool windowed = win != nullptr && win->on;
  uint32_t wox = windowed ? win->ox : 0u, woy = windowed ? win->oy : 0u, woz = windowed ? win->oz : 0u;
  int n_passes = windowed ? win->n_passes : a.n_passes;
  const SortPass* passes = windowed ? win->pass : a.pass;
  // pass p writes buffer pair (p + first) & 1; `first` makes the last pass land where the host expects it
  int first = windowed ? ((win->final_buf - (n_passes - 1)) & 1) : 0;
  uint32_t* violation = windowed ? win->violation : nullptr;
  const int rel_bits = windowed ? win->rel_bits : 32;
  uint32_t* wh = smem;                                   // [kSortWarps][radix]
  uint32_t* sbase = smem + (size_t)kSortWarps * kMaxRadix;  // [radix] digit bases
  __shared__ uint32_t s_part[kSortWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t* src_k = a.keys_in;
  const uint32_t* src_v = a.vals_in;
  for (int p = 0; p < n_passes; ++p) {
    const int shift = passes[p].shift;
    const uint32_t radix = 1u << passes[p].bits, mask = radix - 1u;
    uint32_t* dst_k = a.keys[(p + first) & 1];
    uint32_t* dst_v = a.vals[(p + first) & 1];
    const bool remap = windowed && p == 0;  // pass 0 reads the absolute keys and carries relative ones on
    // ---- hist: hist[tile][d] = number of keys of the tile with digit d
    for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
      for (uint32_t d = threadIdx.x; d < radix; d += kSortThreads) wh[d] = 0;
      __syncthreads();
      const uint32_t wbase = tile * (kSortWarps * 32u * items) + warp * (32u * items);
      uint32_t hk[kSortItemsPerLane];
#pragma unroll
      for (int r = 0; r < kSortItemsPerLane; ++r) {  // all loads first: one memory round trip
        if (r >= (int)items) break;
        const uint32_t i = wbase + r * 32 + lane;
        hk[r] = i < a.n ? __ldcg(src_k + i) : 0u;
        if (remap) {
          hk[r] = window_key(hk[r], wox, woy, woz);
          if (violation && i < a.n && (hk[r] >> rel_bits)) atomicOr(violation, 1u);  // outside the predicted window
        }
      }
#pragma unroll
      for (int r = 0; r < kSortItemsPerLane; ++r) {
        if (r >= (int)items) break;
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < a.n;
        const uint32_t d = valid ? ((hk[r] >> shift) & mask) : (radix + (uint32_t)lane);
        const uint32_t peers = __match_any_sync(kFullMask, d);
        if (valid && (peers & lanemask_lt()) == 0) atomicAdd(&wh[d], (uint32_t)__popc(peers));
      }
      __syncthreads();
      for (uint32_t d = threadIdx.x; d < radix; d += kSortThreads) a.hist[(size_t)d * a.n_tiles + tile] = wh[d];
      __syncthreads();
    }
    grid.sync();
    if (violation && p == 0 && __ldcg(violation)) {
      // the predicted window missed a key (the state was replaced, or a particle outran the CFL bound): every
      // thread sees the flag after the barrier and restarts this pass on the absolute keys
      windowed = false;
      win->on = false;
      violation = nullptr;
      wox = woy = woz = 0u;
      n_passes = a.n_passes;
      passes = a.pass;
      first = 0;
      p = -1;
      continue;
    }
    if (win && win->stamp_out && p == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      win->stamp_out[0] = (uint32_t)(t - win->t0);
    }
    // ---- scan: one warp per digit, exclusive over the tiles (in place, rows are contiguous);
    // totals[d] = digit count.  The row is read 8 x 32 tiles at a time before any of it is rewritten.
    {
      const uint32_t gw = (blockIdx.x * kSortThreads + threadIdx.x) >> 5, n_gw = (gridDim.x * kSortThreads) >> 5;
      for (uint32_t d = gw; d < radix; d += n_gw) {
        uint32_t* row = a.hist + (size_t)d * a.n_tiles;
        uint32_t run = 0;
        for (uint32_t t0 = 0; t0 < a.n_tiles; t0 += 256) {
          uint32_t c[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t t = t0 + u * 32 + lane;
            c[u] = t < a.n_tiles ? __ldcg(row + t) : 0u;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t t = t0 + u * 32 + lane;
            uint32_t inc = c[u];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const uint32_t v = __shfl_up_sync(kFullMask, inc, o);
              if (lane >= o) inc += v;
            }
            if (t < a.n_tiles) row[t] = run + inc - c[u];
            run += __shfl_sync(kFullMask, inc, 31);
          }
        }
        if (lane == 0) a.totals[d] = run;
      }
    }
    grid.sync();
    if (win && win->stamp_out && p == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      win->stamp_out[1] = (uint32_t)(t - win->t0);
    }
    // ---- scatter
    if (blockIdx.x < a.n_tiles) {
      // digit bases: exclusive scan of totals[] (every CTA computes its own copy)
      const uint32_t per = radix / kSortThreads ? radix / kSortThreads : 1u;  // digits per thread (radix >= 256) or 1
      const uint32_t d0 = threadIdx.x * per;
      uint32_t loc[kMaxRadix / kSortThreads];
      uint32_t sum = 0;
#pragma unroll
      for (uint32_t e = 0; e < kMaxRadix / kSortThreads; ++e) {
        loc[e] = (e < per && d0 + e < radix) ? a.totals[d0 + e] : 0u;
        sum += loc[e];
      }
      uint32_t inc = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(kFullMask, inc, o);
        if (lane >= o) inc += v;
      }
      if (lane == 31) s_part[warp] = inc;
      __syncthreads();
      uint32_t wofs = 0;
      for (int w = 0; w < warp; ++w) wofs += s_part[w];
      uint32_t run = wofs + inc - sum;
#pragma unroll
      for (uint32_t e = 0; e < kMaxRadix / kSortThreads; ++e) {
        if (e < per && d0 + e < radix) sbase[d0 + e] = run;
        run += loc[e];
      }
      __syncthreads();
    }
    for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
      {
        uint4* z = reinterpret_cast<uint4*>(wh);  // radix * kSortWarps is a multiple of 4
        for (uint32_t d = threadIdx.x; d < radix * kSortWarps / 4; d += kSortThreads) z[d] = make_uint4(0u, 0u, 0u, 0u);
      }
      __syncthreads();
      const uint32_t wbase = tile * (kSortWarps * 32u * items) + warp * (32u * items);
      uint32_t key[kSortItemsPerLane], val[kSortItemsPerLane];
      uint32_t* mywh = wh + (size_t)warp * radix;
#pragma unroll
      for (int r = 0; r < kSortItemsPerLane; ++r) {  // all loads first: one memory round trip
        if (r >= (int)items) break;
        const uint32_t i = wbase + r * 32 + lane;
        key[r] = i < a.n ? __ldcg(src_k + i) : 0u;
        if (remap) key[r] = window_key(key[r], wox, woy, woz);
        val[r] = (src_v && i < a.n) ? __ldcg(src_v + i) : i;
      }
      // this tile's digit offsets (one more round trip, overlapped with phase 1)
      uint32_t tofs[kMaxRadix / kSortThreads];
#pragma unroll
      for (uint32_t e = 0; e < kMaxRadix / kSortThreads; ++e) {
        const uint32_t d = threadIdx.x + e * kSortThreads;
        tofs[e] = d < radix ? __ldcg(a.hist + (size_t)d * a.n_tiles + tile) : 0u;
      }
      // phase 1: per-warp digit histogram (warp w owns the items [w*512, (w+1)*512) of the tile)
#pragma unroll
      for (int r = 0; r < kSortItemsPerLane; ++r) {
        if (r >= (int)items) break;
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < a.n;
        const uint32_t d = valid ? ((key[r] >> shift) & mask) : (radix + (uint32_t)lane);
        const uint32_t peers = __match_any_sync(kFullMask, d);
        if (valid && (peers & lanemask_lt()) == 0) mywh[d] += __popc(peers);
        __syncwarp();
      }
      __syncthreads();
      // phase 2: per-warp counts of digit d -> global start offsets
#pragma unroll
      for (uint32_t e = 0; e < kMaxRadix / kSortThreads; ++e) {
        const uint32_t d = threadIdx.x + e * kSortThreads;
        if (d >= radix) break;
        uint32_t run = tofs[e] + sbase[d];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
          const uint32_t c = wh[(size_t)w * radix + d];
          wh[(size_t)w * radix + d] = run;
          run += c;
        }
      }
      __syncthreads();
      // phase 3: same walk, rank inside the peer group, scatter
#pragma unroll
      for (int r = 0; r < kSortItemsPerLane; ++r) {
        if (r >= (int)items) break;
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < a.n;
        const uint32_t d = valid ? ((key[r] >> shift) & mask) : (radix + (uint32_t)lane);
        const uint32_t peers = __match_any_sync(kFullMask, d);
        const uint32_t rank = __popc(peers & lanemask_lt());
        uint32_t dst = 0;
        if (valid) dst = mywh[d] + rank;
        __syncwarp();
        if (valid && rank == 0) mywh[d] += __popc(peers);
        __syncwarp();
        if (valid) {
          dst_k[dst] = key[r];
          dst_v[dst] = val[r];
        }
      }
      __syncthreads();
    }
    grid.sync();
    if (win && win->stamp_out && p == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      win->stamp_out[2] = (uint32_t)(t - win->t0);
    }
    src_k = dst_k;
    src_v = dst_v;
  }
  return (n_passes - 1 + first) & 1;
}

__global__ void __launch_bounds__(kSortThreads, 2) radix_sort_kernel(SortArgs a) {
  extern __shared__ __align__(16) uint32_t sort_smem[];
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  radix_sort_device(a, sort_smem, grid);
}

// Host helper: fills the pass plan and buffers of a SortArgs.
inline SortArgs make_sort_args(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys[2], uint32_t* vals[2],
                               uint32_t n, int num_bit, uint32_t* scratch) {
  SortArgs a{};
  a.keys_in = keys_in;
  a.vals_in = vals_in;
  a.keys[0] = keys[0];
  a.keys[1] = keys[1];
  a.vals[0] = vals[0];
  a.vals[1] = vals[1];
  a.n = n;
  a.items = sort_items_for(n);
  a.n_tiles = (n + kSortWarps * 32u * a.items - 1) / (kSortWarps * 32u * a.items);
  a.n_passes = plan_sort_passes(num_bit, a.pass);
  a.hist = scratch;
  a.totals = scratch + (size_t)kMaxRadix * (a.n_tiles ? a.n_tiles : 1);
  return a;
}

// Cooperative-launch geometry of a kernel that embeds radix_sort_device.
template <typename Kernel>
inline int sort_grid_ctas(Kernel kernel, int sm_count, uint32_t n_tiles) {
  int occ = 0;
  GMPM_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSortSmemBytes));
  GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kSortThreads, kSortSmemBytes));
  if (occ < 1) throw CudaError(cudaErrorLaunchOutOfResources, "radix sort kernel does not fit on an SM");
  const uint32_t max_ctas = (uint32_t)(sm_count * occ);
  const uint32_t want = n_tiles > 64u ? n_tiles : 64u;  // the scan phase wants a few warps even for tiny inputs
  return (int)(want < max_ctas ? want : max_ctas);
}

}  // namespace gmpm
