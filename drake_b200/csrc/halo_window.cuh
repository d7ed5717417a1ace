// halo_window.cuh -- layout of and access to the peer-memory receive windows of the shared-block exchange
// (protocol: halo.cuh).  Separate from halo.cuh because the grid kernels (kernels.cuh) embed the send / receive
// halves when the exchange is fused into the grid assembly and update.
#pragma once
#include "common.cuh"

namespace gmpm {

constexpr int kHaloMaxPeers = 8;
constexpr int kHaloPeerThreads = 128;                              // one CTA per listed block, two values per thread
constexpr unsigned long long kHaloPeerTimeoutNs = 20000000000ull;  // 20 s; a delivery takes microseconds
constexpr uint32_t kHaloPeerTimeoutCode = 1u << 16;               // added to the error word the next exchange checks
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
struct HaloWindows {
  unsigned char* base[kHaloMaxPeers];
  // false (default): every thread that stored into a peer window fences at SYSTEM scope itself -- the documented
  // requirement for peer visibility.  true (GMPM_HALO_GPU_SCOPE_FENCE=1): the storers fence at GPU scope and only the
  // last arriver fences at system scope (valid under PTX cross-scope cumulativity; ~8 us faster per substep; kept
  // behind the switch until the multi-rank stress test has run on it, tests/test_slab.py).
  uint32_t gpu_scope_fence;
};
// Window layout: flag[64] | present[2][world][cap] | (256-byte aligned) data[2][world][cap][256]
__host__ __device__ inline size_t halo_window_data_offset(uint32_t world, uint32_t cap) {
  return (256 + sizeof(uint32_t) * 2 * (size_t)world * cap + 255) & ~(size_t)255;
}
template <typename T>
__host__ __device__ inline size_t halo_window_bytes(uint32_t world, uint32_t cap) {
  return halo_window_data_offset(world, cap) + sizeof(T) * 256 * 2 * (size_t)world * cap;
}
__device__ __forceinline__ uint32_t* halo_window_present(unsigned char* base, uint32_t parity, uint32_t world,
                                                         uint32_t src, uint32_t cap, uint32_t i) {
  return reinterpret_cast<uint32_t*>(base + 256) + ((size_t)parity * world + src) * cap + i;
}
template <typename T>
__device__ __forceinline__ T* halo_window_data(unsigned char* base, uint32_t parity, uint32_t world, uint32_t src,
                                               uint32_t cap, uint32_t i) {
  return reinterpret_cast<T*>(base + halo_window_data_offset(world, cap)) + (((size_t)parity * world + src) * cap + i) * 256;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
template <typename T> struct Pair;
template <> struct Pair<float> { using type = float2; };
template <> struct Pair<double> { using type = double2; };
// the 256 values of a block, [mass 64 | momentum 3 x 64], as 128 pairs: pair t of block g
template <typename T>
__device__ __forceinline__ typename Pair<T>::type* halo_block_pair(T* g_mass, T* g_vel, uint32_t g, uint32_t t) {
  T* p = t < 32u ? g_mass + (size_t)g * 64 + 2u * t : g_vel + (size_t)g * 192 + 2u * (t - 32u);
  return reinterpret_cast<typename Pair<T>::type*>(p);
}


constexpr uint32_t kHaloNoIndex = 0xFFFFFFFFu;
// What a grid kernel needs to take part in the exchange (HALO = 1: send from the assembly, 2: receive in the update).
template <typename T>
struct HaloFuse {
  const uint32_t* index;     // grid block -> position in the shared-block list of this exchange, or kHaloNoIndex
  const uint32_t* list;      // the shared-block list (ascending block ids) and its length
  uint32_t n_list;
  const uint32_t* touched_epoch;  // touched_epoch[g] == p2g_epoch: this rank's P2G reached block g
  uint32_t p2g_epoch;
  const uint8_t* sharers;    // sharer set per list position
  HaloWindows win;           // every rank's window
  uint32_t rank, world, cap, xchg;
  const uint32_t* peers;     // ranks this rank shares a block with (device word)
  uint32_t* done;            // CTA arrival counter of the send
  uint32_t* err;             // error word (timeouts)
};

// Waits (lanes < world of the calling warp) until every OTHER sharer in `who` has delivered exchange `xchg`.
__device__ __forceinline__ void halo_wait_sharers(const unsigned char* window, uint32_t who, uint32_t rank,
                                                  uint32_t world, uint32_t xchg, uint32_t* err) {
  const uint32_t lane = threadIdx.x & 31u;
  if (lane < world && lane != rank && ((who >> lane) & 1u)) {
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(window) + lane;
    const unsigned long long t0 = global_timer_ns();
    while ((int32_t)(ld_acquire_sys(flag) - xchg) < 0) {
      __nanosleep(32);
      if (global_timer_ns() - t0 > kHaloPeerTimeoutNs) {  // a peer died or the call order differs between ranks
        atomicAdd(err, kHaloPeerTimeoutCode);
        break;
      }
    }
  }
  __syncwarp();
}

// End of a sending kernel: after the stores of ALL its CTAs (threadFenceReduction pattern) the last CTA raises
// flag[rank] = xchg in the window of every peer.  Call with all threads of the CTA.
// Scopes (win.gpu_scope_fence = 1 only; by default every storer fences at system scope): the storing threads fence
// at GPU scope only -- a system-scope fence per warp costs ~20 us per kernel even
// with purely local stores (measured with two states on one GPU, tools/probe_exchange.py: they serialise) -- and
// arrive on a GPU-scope counter; the single last arriver then fences at SYSTEM scope before its release stores.
// PTX causality order is cumulative across scopes: store -> fence.gpu -> arrive  synchronises with  last arrive ->
// fence.sys -> st.release.sys flag  synchronises with the receiver's ld.acquire.sys, so the data is visible there.
__device__ __forceinline__ void halo_signal_peers(const HaloWindows& win, uint32_t rank, uint32_t world, uint32_t xchg,
                                                  const uint32_t* peers_mask, uint32_t* done) {
  if (win.gpu_scope_fence) __threadfence(); else __threadfence_system();
  __syncthreads();
  __shared__ bool s_halo_last;
  if (threadIdx.x == 0) s_halo_last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_halo_last) return;
  if (threadIdx.x == 0) *done = 0;
  const uint32_t peers = *peers_mask & ~(1u << rank);
  if (threadIdx.x < world && ((peers >> threadIdx.x) & 1u)) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(win.base[threadIdx.x]) + rank, xchg);
  }
}

// The same per WARP and without a CTA barrier, for a sender that has other work to do afterwards (the grid
// assembly): a warp that stored to a peer fences, every warp of the grid arrives once, the last one raises the flags.
// Warps that sent nothing pass straight through; the fence of the others is hidden by the rest of the SM.
__device__ __forceinline__ void halo_signal_peers_warp(const HaloWindows& win, uint32_t rank, uint32_t world,
                                                       uint32_t xchg, const uint32_t* peers_mask, uint32_t* done,
                                                       bool sent) {
  if (__any_sync(0xFFFFFFFFu, sent)) {
    if (win.gpu_scope_fence) __threadfence();  // GPU scope; the last arriver fences at system scope
    else __threadfence_system();
  }
  __syncwarp();
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t last = 0;
  if (lane == 0) last = atomicAdd(done, 1u) == gridDim.x * (blockDim.x >> 5) - 1u ? 1u : 0u;
  last = __shfl_sync(0xFFFFFFFFu, last, 0);
  if (!last) return;
  if (lane == 0) *done = 0;
  const uint32_t peers = *peers_mask & ~(1u << rank);
  if (lane < world && ((peers >> lane) & 1u)) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(win.base[lane]) + rank, xchg);
  }
}

}  // namespace gmpm
