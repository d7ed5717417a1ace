// sort.cuh -- stable LSD radix sort of (key, value) uint32 pairs on key bits [0, num_bit).
//
// Replaces the reference's radix_sort (radix_sort.cuh:202-286: 2-bit digits, Hillis-Steele block
// scans, recursive Blelloch scan, ~5 global sweeps per digit, i.e. 8 digit passes for 16 bits).
// Here: 8-bit digits, so 16 key bits = 2 passes and a full 21-bit cell key = 3 passes; each pass is
//   (1) radix_hist   : per-tile digit histogram  -> hist[digit][tile], digit totals
//   (2) radix_scan   : one CTA per digit: exclusive scan over tiles + digit base
//   (3) radix_scatter: stable ranking inside the tile with warp match-any, scatter to global
// Stability (ties keep input order) is what makes the result bit-identical to the reference's
// permutation.  Traffic per pass: 4 B (hist) + 8 B read + 8 B write per element.
#pragma once
#include "common.cuh"

namespace gmpm {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItemsPerLane = 16;                                  // per pass-through of a warp
constexpr int kSortWarpChunk = 32 * kSortItemsPerLane;                 // 512 consecutive items / warp
constexpr int kSortTile = kSortWarps * kSortWarpChunk;                 // 4096 items / CTA

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// hist[d * n_tiles + tile] = number of keys of the tile with digit d; totals[d] += same.
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, int shift, uint32_t digit_mask,
                  uint32_t n_tiles, uint32_t* __restrict__ hist, uint32_t* __restrict__ totals) {
  __shared__ uint32_t sh[kRadix];
  const uint32_t tile = blockIdx.x;
  sh[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t begin = tile * kSortTile;
  const uint32_t end = min(begin + (uint32_t)kSortTile, n);
  for (uint32_t i = begin + threadIdx.x; i < end; i += kSortThreads) {
    const uint32_t d = (keys[i] >> shift) & digit_mask;
    atomicAdd(&sh[d], 1u);
  }
  __syncthreads();
  const uint32_t c = sh[threadIdx.x];
  hist[threadIdx.x * n_tiles + tile] = c;
  if (c) atomicAdd(&totals[threadIdx.x], c);
}

// One CTA per digit d: offsets[d][tile] = (sum of totals of digits < d) + exclusive scan over tiles.
__global__ void __launch_bounds__(kSortThreads)
radix_scan_kernel(uint32_t* __restrict__ hist, const uint32_t* __restrict__ totals, uint32_t n_tiles) {
  __shared__ uint32_t sh[kSortThreads];
  __shared__ uint32_t s_base;
  const uint32_t d = blockIdx.x;
  // digit base
  sh[threadIdx.x] = threadIdx.x < d ? totals[threadIdx.x] : 0u;
  __syncthreads();
  for (int off = kSortThreads / 2; off > 0; off >>= 1) {
    if (threadIdx.x < off) sh[threadIdx.x] += sh[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) s_base = sh[0];
  __syncthreads();
  const uint32_t base = s_base;
  // chunked exclusive scan of the row
  uint32_t* row = hist + (size_t)d * n_tiles;
  const uint32_t per = (n_tiles + kSortThreads - 1) / kSortThreads;
  const uint32_t b = min(threadIdx.x * per, n_tiles), e = min(b + per, n_tiles);
  uint32_t sum = 0;
  for (uint32_t i = b; i < e; ++i) sum += row[i];
  __syncthreads();
  sh[threadIdx.x] = sum;
  __syncthreads();
  // Hillis-Steele inclusive scan over 256 partials
  for (int off = 1; off < kSortThreads; off <<= 1) {
    uint32_t v = threadIdx.x >= off ? sh[threadIdx.x - off] : 0u;
    __syncthreads();
    sh[threadIdx.x] += v;
    __syncthreads();
  }
  uint32_t run = base + sh[threadIdx.x] - sum;
  for (uint32_t i = b; i < e; ++i) {
    const uint32_t c = row[i];
    row[i] = run;
    run += c;
  }
}

// Stable scatter.  Warp w of the CTA owns the 512 consecutive items [w*512, (w+1)*512) of the tile,
// walked as 16 rounds of 32 lanes, so input order = (warp, round, lane).
template <bool IOTA_VALUES>
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
                     int shift, uint32_t digit_mask, uint32_t n_tiles,
                     const uint32_t* __restrict__ offsets) {
  __shared__ uint32_t wh[kSortWarps][kRadix];
  const uint32_t tile = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int w = 0; w < kSortWarps; ++w) wh[w][threadIdx.x] = 0;
  __syncthreads();
  const uint32_t wbase = tile * kSortTile + warp * kSortWarpChunk;
  uint32_t key[kSortItemsPerLane];
  // phase 1: per-warp digit histogram
#pragma unroll
  for (int r = 0; r < kSortItemsPerLane; ++r) {
    const uint32_t i = wbase + r * 32 + lane;
    const bool valid = i < n;
    key[r] = valid ? keys_in[i] : 0u;
    const uint32_t d = valid ? ((key[r] >> shift) & digit_mask) : (uint32_t)(kRadix + lane);
    const uint32_t peers = __match_any_sync(kFullMask, d);
    if (valid && (peers & lanemask_lt()) == 0) wh[warp][d] += __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  // phase 2: thread d turns the per-warp counts of digit d into global start offsets
  {
    const uint32_t d = threadIdx.x;
    uint32_t run = offsets[d * n_tiles + tile];
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
      const uint32_t c = wh[w][d];
      wh[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
  // phase 3: same walk, rank inside the peer group, scatter
#pragma unroll
  for (int r = 0; r < kSortItemsPerLane; ++r) {
    const uint32_t i = wbase + r * 32 + lane;
    const bool valid = i < n;
    const uint32_t d = valid ? ((key[r] >> shift) & digit_mask) : (uint32_t)(kRadix + lane);
    const uint32_t peers = __match_any_sync(kFullMask, d);
    const uint32_t rank = __popc(peers & lanemask_lt());
    uint32_t dst = 0;
    if (valid) dst = wh[warp][d] + rank;
    __syncwarp();
    if (valid && rank == 0) wh[warp][d] += __popc(peers);
    __syncwarp();
    if (valid) {
      keys_out[dst] = key[r];
      vals_out[dst] = IOTA_VALUES ? i : vals_in[i];
    }
  }
}

// Scratch needed by radix_sort_pairs for n elements (uint32 words).
inline size_t radix_sort_scratch_words(size_t n) {
  const size_t n_tiles = (n + kSortTile - 1) / kSortTile;
  return (size_t)kRadix * (n_tiles ? n_tiles : 1) + 4 * kRadix;
}

struct LaunchCounter;  // fwd (gmpm.cu)

// Sorts (keys[0], vals[0]) using (keys[1], vals[1]) as the ping-pong partner.  Returns the index
// (0 or 1) of the buffer pair holding the result.  If iota_first, the values of the first pass are
// generated (0..n-1) instead of read.  `on_launch(name)` is called once per kernel launch.
template <typename F>
inline int radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], uint32_t n, int num_bit, bool iota_first,
                            uint32_t* scratch, cudaStream_t stream, F&& on_launch) {
  if (n == 0 || num_bit <= 0) {
    return 0;
  }
  const uint32_t n_tiles = (n + kSortTile - 1) / kSortTile;
  uint32_t* hist = scratch;
  uint32_t* totals = scratch + (size_t)kRadix * n_tiles;
  const int passes = (num_bit + kRadixBits - 1) / kRadixBits;
  GMPM_CUDA(cudaMemsetAsync(totals, 0, sizeof(uint32_t) * kRadix * 4, stream));
  int cur = 0;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * kRadixBits;
    const int bits = (num_bit - shift) < kRadixBits ? (num_bit - shift) : kRadixBits;
    const uint32_t mask = (1u << bits) - 1u;
    uint32_t* tot = totals + (size_t)(p & 3) * kRadix;
    if (p >= 4) GMPM_CUDA(cudaMemsetAsync(tot, 0, sizeof(uint32_t) * kRadix, stream));
    on_launch("radix_hist");
    radix_hist_kernel<<<n_tiles, kSortThreads, 0, stream>>>(keys[cur], n, shift, mask, n_tiles, hist, tot);
    on_launch("radix_scan");
    radix_scan_kernel<<<kRadix, kSortThreads, 0, stream>>>(hist, tot, n_tiles);
    on_launch("radix_scatter");
    if (p == 0 && iota_first)
      radix_scatter_kernel<true><<<n_tiles, kSortThreads, 0, stream>>>(keys[cur], vals[cur], keys[cur ^ 1],
                                                                       vals[cur ^ 1], n, shift, mask, n_tiles, hist);
    else
      radix_scatter_kernel<false><<<n_tiles, kSortThreads, 0, stream>>>(keys[cur], vals[cur], keys[cur ^ 1],
                                                                        vals[cur ^ 1], n, shift, mask, n_tiles, hist);
    GMPM_CUDA(cudaGetLastError());
    cur ^= 1;
  }
  return cur;
}

}  // namespace gmpm
