/*
 * gmpm_oracle.c -- CPU restatement of g1n0st/drake multibody/gpu_mpm.  TEST INFRASTRUCTURE ONLY
 * (see gmpm_oracle.h).  Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC.
 */
#include "gmpm_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ref: settings.h:37-126 */
void ogmpm_default_params(ogmpm_params* p) {
  p->domain_bits = 7;
  p->boundary_cells = 3;
  p->gravity_axis = 2;
  p->reserved = 0;
  p->youngs = 400000.;
  p->poisson = .3;
  p->density = 2000.;
  p->gamma = 0.;
  p->k_normal = 100000.;
  p->v_blend = .8;
  p->c_f = 0.;
  p->sdf_friction = 0.3;
  p->gravity = -9.8;
  p->epsv = 1e-3;
  p->dx = 0.;
}

/* ref: cuda_mpm_kernels.cuh:296-304 */
uint32_t ogmpm_contract_bits(uint32_t v) {
  v &= 0x09249249u;
  v = (v ^ (v >> 2)) & 0x030C30C3u;
  v = (v ^ (v >> 4)) & 0x0300F00Fu;
  v = (v ^ (v >> 8)) & 0xFF0000FFu;
  v = (v ^ (v >> 16)) & 0x000003FFu;
  return v;
}

/* ref: cuda_mpm_kernels.cuh:306-313 */
uint32_t ogmpm_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

/* ref: cuda_mpm_kernels.cuh:317-323 (x is the most significant of each bit triple) */
uint32_t ogmpm_morton_code(uint32_t x, uint32_t y, uint32_t z) {
  return ogmpm_expand_bits(x) * 4u + ogmpm_expand_bits(y) * 2u + ogmpm_expand_bits(z);
}

/* ref: cuda_mpm_kernels.cuh:334-341 (BLOCK_BITS = 2, settings.h:48) */
uint32_t ogmpm_cell_index(uint32_t xi, uint32_t yi, uint32_t zi) {
  uint32_t higher_bit = ogmpm_morton_code(xi >> 2, yi >> 2, zi >> 2);
  uint32_t lower_bit = ((xi & 3u) << 4) | ((yi & 3u) << 2) | (zi & 3u);
  return (higher_bit << 6) | lower_bit;
}

/* ref: cuda_mpm_kernels.cuh:325-332, 343-363 */
void ogmpm_inverse_cell_index(uint32_t index, uint32_t* xyz) {
  uint32_t higher_bit = index >> 6;
  uint32_t lower_bit = index & 63u;
  uint32_t lx = (lower_bit >> 4) & 3u, ly = (lower_bit >> 2) & 3u, lz = lower_bit & 3u;
  uint32_t hx = ogmpm_contract_bits(higher_bit >> 2);
  uint32_t hy = ogmpm_contract_bits(higher_bit >> 1);
  uint32_t hz = ogmpm_contract_bits(higher_bit);
  xyz[0] = (hx << 2) | lx;
  xyz[1] = (hy << 2) | ly;
  xyz[2] = (hz << 2) | lz;
}

/* ref: radix_sort.cuh:202-286.  The reference runs num_bit/2 stable 2-bit LSD passes; the net
 * effect is a stable sort by (key & ((1 << num_bit) - 1)), restated as a stable counting sort per
 * 8-bit digit.  Keys keep all their bits; ties keep input order. */
void ogmpm_stable_sort_low_bits(const uint32_t* keys_in, const uint32_t* ids_in, size_t n, int num_bit,
                                uint32_t* keys_out, uint32_t* ids_out) {
  uint32_t* ka = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
  uint32_t* va = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
  uint32_t* kb = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
  uint32_t* vb = (uint32_t*)malloc((n ? n : 1) * sizeof(uint32_t));
  memcpy(ka, keys_in, n * sizeof(uint32_t));
  memcpy(va, ids_in, n * sizeof(uint32_t));
  for (int shift = 0; shift < num_bit; shift += 2) { /* 2-bit digits exactly as the reference */
    size_t cnt[4] = {0, 0, 0, 0}, off[4];
    for (size_t i = 0; i < n; ++i) cnt[(ka[i] >> shift) & 3u]++;
    off[0] = 0;
    for (int d = 1; d < 4; ++d) off[d] = off[d - 1] + cnt[d - 1];
    for (size_t i = 0; i < n; ++i) {
      size_t d = (ka[i] >> shift) & 3u;
      kb[off[d]] = ka[i];
      vb[off[d]] = va[i];
      off[d]++;
    }
    uint32_t* t = ka; ka = kb; kb = t;
    t = va; va = vb; vb = t;
  }
  memcpy(keys_out, ka, n * sizeof(uint32_t));
  memcpy(ids_out, va, n * sizeof(uint32_t));
  free(ka); free(va); free(kb); free(vb);
}

#define T float
#define SUF f32
#define SQRT sqrtf
#define FABS fabsf
#define LOG logf
#define EXP expf
#include "gmpm_oracle_impl.inc"
#undef T
#undef SUF
#undef SQRT
#undef FABS
#undef LOG
#undef EXP

#define T double
#define SUF f64
#define SQRT sqrt
#define FABS fabs
#define LOG log
#define EXP exp
#include "gmpm_oracle_impl.inc"
#undef T
#undef SUF
#undef SQRT
#undef FABS
#undef LOG
#undef EXP
