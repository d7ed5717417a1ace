// common.cuh -- shared definitions of libgmpm_b200 (sm_100a).  Product code: no CPU fallback.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

namespace gmpm {

struct CudaError : std::runtime_error {
  cudaError_t code;
  CudaError(cudaError_t c, const std::string& what) : std::runtime_error(what), code(c) {}
};

inline void cuda_check(cudaError_t e, const char* expr, const char* file, int line) {
  if (e != cudaSuccess) {
    char buf[512];
    snprintf(buf, sizeof(buf), "%s:%d: %s -> %s (%s)", file, line, expr, cudaGetErrorName(e),
             cudaGetErrorString(e));
    throw CudaError(e, buf);
  }
}
#define GMPM_CUDA(expr) ::gmpm::cuda_check((expr), #expr, __FILE__, __LINE__)

constexpr int kBlockBits = 2;                      // settings.h:48  4^3-cell blocks
constexpr int kBlockCells = 64;                    // settings.h:66
constexpr int kTileDim = 6;                        // block + the two-cell reach of the quadratic stencil
constexpr int kTileNodes = kTileDim * kTileDim * kTileDim;  // 216
constexpr unsigned kFullMask = 0xFFFFFFFFu;

// ---- Morton / block-major cell addressing (reference: cuda_mpm_kernels.cuh:296-363) -------------
// key = Morton3(x>>2, y>>2, z>>2) << 6 | (x&3)<<4 | (y&3)<<2 | (z&3), x most significant in Morton.
__host__ __device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__host__ __device__ __forceinline__ uint32_t contract_bits(uint32_t v) {
  v &= 0x09249249u;
  v = (v ^ (v >> 2)) & 0x030C30C3u;
  v = (v ^ (v >> 4)) & 0x0300F00Fu;
  v = (v ^ (v >> 8)) & 0xFF0000FFu;
  v = (v ^ (v >> 16)) & 0x000003FFu;
  return v;
}
__host__ __device__ __forceinline__ uint32_t morton3(uint32_t x, uint32_t y, uint32_t z) {
  return expand_bits(x) * 4u + expand_bits(y) * 2u + expand_bits(z);
}
__host__ __device__ __forceinline__ uint32_t cell_key(uint32_t xi, uint32_t yi, uint32_t zi) {
  return (morton3(xi >> 2, yi >> 2, zi >> 2) << 6) | ((xi & 3u) << 4) | ((yi & 3u) << 2) | (zi & 3u);
}
__host__ __device__ __forceinline__ void block_coords(uint32_t block, uint32_t& bx, uint32_t& by,
                                                      uint32_t& bz) {
  bx = contract_bits(block >> 2);
  by = contract_bits(block >> 1);
  bz = contract_bits(block);
}
__host__ __device__ __forceinline__ void inverse_cell_key(uint32_t key, uint32_t& x, uint32_t& y,
                                                          uint32_t& z) {
  uint32_t bx, by, bz;
  block_coords(key >> 6, bx, by, bz);
  x = (bx << 2) | ((key >> 4) & 3u);
  y = (by << 2) | ((key >> 2) & 3u);
  z = (bz << 2) | (key & 3u);
}

// Device constants of one state (runtime copy of settings.h).  Passed by value to kernels.
template <typename T>
struct Consts {
  int domain_bits;
  int domain_size;      // cells per axis
  uint32_t n_cells;
  uint32_t n_blocks;
  int boundary_cells;
  int gravity_axis;
  T dx, dx_inv, d_inv;  // settings.h:57-59
  T mu, lambda;         // Lame, settings.h:114-115
  T density, gamma, k_normal, v_blend, c_f, sdf_friction, gravity, epsv;
};

// Base node of the quadratic stencil (cuda_mpm_kernels.cuh:365-382).  The float->uint32 conversion
// saturates at 0 for negative input (cvt.rzi.u32), as in the reference; additionally clamped so
// that the 3-wide stencil never leaves the grid (memory safety; never active for in-domain scenes).
template <typename T>
__device__ __forceinline__ uint32_t base_node(T x, T dx_inv, int domain_size) {
  T v = x * dx_inv - T(0.5);
  uint32_t b = (v > T(0)) ? static_cast<uint32_t>(v) : 0u;
  uint32_t hi = static_cast<uint32_t>(domain_size - 3);
  return b < hi ? b : hi;
}

}  // namespace gmpm
