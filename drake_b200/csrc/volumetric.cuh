// volumetric.cuh -- particle-only (volumetric) MPM materials (absent upstream, parity UNPINNED: the
// checker is oracle/gmpm_oracle_vol.inc, a CPU restatement of the same published algorithms).
//
// The reference has one entry point (AddQRCloth, cuda_mpm_model.cu:16-33) and one constitutive law (the QR cloth,
// cuda_mpm_kernels.cuh:72-181).  This file generalises that slot of the substep -- CalcFemStateAndForce -- to
// particles that carry their own 3x3 deformation gradient:
//     F <- (I + dt C) F                      C = grad v from the previous G2P (MLS-MPM, Hu et al. 2018, eq. 17)
//     F = U Sigma V^T                        one-sided Jacobi on F^T F, U and V rotations, sign in sigma_3
//     fixed corotated   (Stomakhin et al. 2013, eq. 1):  tau_i = 2 mu (s_i - 1) s_i + lambda (J - 1) J
//     StVK-Hencky       (Klar et al. 2016, eq. 26):      tau_i = 2 mu e_i + lambda tr(e),  e = log Sigma
//     Drucker-Prager    (Klar et al. 2016, sec. 6.2 / alg. 2): return mapping of e onto the cone
//                        |dev e| + alpha (3 lambda + 2 mu) / (2 mu) tr(e) <= 0, tip at tr(e) >= 0; F <- U exp(e) V^T
//     tau = U diag(tau_i) U^T (Kirchhoff), stored as V0 tau in the 6 scalars the P2G kernels read
//     (xx, yy, zz, xy, xz, yz): P2G adds -dt D^-1 V0 tau to m C exactly as for the cloth faces (:438-446).
#pragma once
#include "../../include/gmpm_c.h"
#include "common.cuh"
#include "math.cuh"

namespace gmpm {

template <typename T> __device__ __forceinline__ T tlog(T x);
template <> __device__ __forceinline__ float tlog<float>(float x) { return logf(x); }
template <> __device__ __forceinline__ double tlog<double>(double x) { return log(x); }
template <typename T> __device__ __forceinline__ T texp(T x);
template <> __device__ __forceinline__ float texp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ double texp<double>(double x) { return exp(x); }

constexpr int kMaxMaterials = 16;
constexpr int kSvdSweeps = 8;   // cyclic Jacobi sweeps on the symmetric 3x3 (converged to rounding in <= 6 for fp64)

template <typename T>
struct MaterialDev {
  int model;          // gmpm_material_model
  T mu, lambda;       // Lame parameters
  T alpha;            // Drucker-Prager: sqrt(2/3) 2 sin(phi) / (3 - sin(phi))
  T cohesion;         // Drucker-Prager: shift of the cone tip along tr(e)
};
template <typename T>
struct MaterialTable {
  MaterialDev<T> m[kMaxMaterials];
};

// One Jacobi rotation in the (P, Q) plane of the symmetric A, accumulated into V (columns = eigenvectors).
template <typename T, int P, int Q>
__device__ __forceinline__ void jacobi_rotate(T (&A)[3][3], T (&V)[3][3]) {
  const T apq = A[P][Q];
  if (apq == T(0)) return;
  const T theta = (A[Q][Q] - A[P][P]) / (T(2) * apq);
  const T t = (theta >= T(0) ? T(1) : T(-1)) / (tabs(theta) + tsqrt(theta * theta + T(1)));
  const T c = T(1) / tsqrt(t * t + T(1)), s = t * c;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const T akp = A[k][P], akq = A[k][Q];
    A[k][P] = c * akp - s * akq;
    A[k][Q] = s * akp + c * akq;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const T apk = A[P][k], aqk = A[Q][k];
    A[P][k] = c * apk - s * aqk;
    A[Q][k] = s * apk + c * aqk;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const T vkp = V[k][P], vkq = V[k][Q];
    V[k][P] = c * vkp - s * vkq;
    V[k][Q] = s * vkp + c * vkq;
  }
}

// F = U diag(sig) V^T with U, V proper rotations, sig[0] >= sig[1] >= |sig[2]|, sign of det F in sig[2].
template <typename T>
__device__ inline void svd3(const T* F, T (&U)[3][3], T (&sig)[3], T (&V)[3][3]) {
  T A[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      A[i][j] = F[0 * 3 + i] * F[0 * 3 + j] + F[1 * 3 + i] * F[1 * 3 + j] + F[2 * 3 + i] * F[2 * 3 + j];
      V[i][j] = i == j ? T(1) : T(0);
    }
  // cyclic sweeps until the off-diagonal part is below rounding (quadratic convergence: 3-5 sweeps for the
  // deformation gradients of a simulation; kSvdSweeps bounds the loop)
  const T thr = sizeof(T) == 8 ? T(1e-16) : T(1e-7);
  for (int sweep = 0; sweep < kSvdSweeps; ++sweep) {
    if (tabs(A[0][1]) + tabs(A[0][2]) + tabs(A[1][2]) <= thr * (tabs(A[0][0]) + tabs(A[1][1]) + tabs(A[2][2]))) break;
    jacobi_rotate<T, 0, 1>(A, V);
    jacobi_rotate<T, 0, 2>(A, V);
    jacobi_rotate<T, 1, 2>(A, V);
  }
  T e[3] = {A[0][0], A[1][1], A[2][2]};
  // sort the eigenvalues descending (column swaps of V; a swap flips the handedness, restored below)
#define GMPM_SWAP_COL(a, b)                                                \
  if (e[a] < e[b]) {                                                       \
    const T te = e[a]; e[a] = e[b]; e[b] = te;                             \
    for (int k = 0; k < 3; ++k) { const T tv = V[k][a]; V[k][a] = V[k][b]; V[k][b] = tv; } \
  }
  GMPM_SWAP_COL(0, 1)
  GMPM_SWAP_COL(0, 2)
  GMPM_SWAP_COL(1, 2)
#undef GMPM_SWAP_COL
  const T detV = V[0][0] * (V[1][1] * V[2][2] - V[1][2] * V[2][1]) - V[0][1] * (V[1][0] * V[2][2] - V[1][2] * V[2][0]) +
                 V[0][2] * (V[1][0] * V[2][1] - V[1][1] * V[2][0]);
  if (detV < T(0))
    for (int k = 0; k < 3; ++k) V[k][2] = -V[k][2];
  // B = F V; u0 = b0 / |b0|, u1 = b1 orthogonalised against u0, u2 = u0 x u1, sig2 = u2 . b2 (signed)
  T B[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) B[i][j] = F[i * 3 + 0] * V[0][j] + F[i * 3 + 1] * V[1][j] + F[i * 3 + 2] * V[2][j];
  const T tiny = T(1e-30);
  T n0 = tsqrt(B[0][0] * B[0][0] + B[1][0] * B[1][0] + B[2][0] * B[2][0]);
  T u0[3] = {T(1), T(0), T(0)};
  if (n0 > tiny) { u0[0] = B[0][0] / n0; u0[1] = B[1][0] / n0; u0[2] = B[2][0] / n0; }
  const T d01 = u0[0] * B[0][1] + u0[1] * B[1][1] + u0[2] * B[2][1];
  T r1[3] = {B[0][1] - d01 * u0[0], B[1][1] - d01 * u0[1], B[2][1] - d01 * u0[2]};
  T n1 = tsqrt(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
  T u1[3];
  if (n1 > tiny) {
    u1[0] = r1[0] / n1; u1[1] = r1[1] / n1; u1[2] = r1[2] / n1;
  } else {  // rank <= 1: any unit vector orthogonal to u0
    const T ax = tabs(u0[0]), ay = tabs(u0[1]), az = tabs(u0[2]);
    T a[3] = {T(0), T(0), T(0)};
    if (ax <= ay && ax <= az) a[0] = T(1); else if (ay <= az) a[1] = T(1); else a[2] = T(1);
    const T da = a[0] * u0[0] + a[1] * u0[1] + a[2] * u0[2];
    T q[3] = {a[0] - da * u0[0], a[1] - da * u0[1], a[2] - da * u0[2]};
    const T nq = tsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    u1[0] = q[0] / nq; u1[1] = q[1] / nq; u1[2] = q[2] / nq;
    n1 = T(0);
  }
  const T u2[3] = {u0[1] * u1[2] - u0[2] * u1[1], u0[2] * u1[0] - u0[0] * u1[2], u0[0] * u1[1] - u0[1] * u1[0]};
  sig[0] = n0;
  sig[1] = n1;
  sig[2] = u2[0] * B[0][2] + u2[1] * B[1][2] + u2[2] * B[2][2];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    U[i][0] = u0[i];
    U[i][1] = u1[i];
    U[i][2] = u2[i];
  }
}

// One particle: F (row-major 9, in/out) advanced by grad v, returns the Kirchhoff stress tau (xx, yy, zz, xy, xz, yz).
template <typename T>
__device__ inline void volumetric_update(const MaterialDev<T>& m, const T* C, T dt, T* F, T (&tau)[6]) {
  T Fn[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      Fn[i * 3 + j] = F[i * 3 + j] + dt * (C[i * 3 + 0] * F[0 * 3 + j] + C[i * 3 + 1] * F[1 * 3 + j] + C[i * 3 + 2] * F[2 * 3 + j]);
  T U[3][3], V[3][3], s[3], t[3];
  svd3<T>(Fn, U, s, V);
  if (m.model == GMPM_MAT_FIXED_COROTATED) {
    const T J = s[0] * s[1] * s[2];
    const T vol_term = m.lambda * (J - T(1)) * J;
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = T(2) * m.mu * (s[i] - T(1)) * s[i] + vol_term;
  } else {
    const T floor_s = T(1e-6);
    T e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) e[i] = tlog(s[i] > floor_s ? s[i] : floor_s);
    if (m.model == GMPM_MAT_DRUCKER_PRAGER) {
      // Klar et al. 2016, alg. 2 (no hardening); cohesion shifts the tip
      const T tr = e[0] + e[1] + e[2] - T(3) * m.cohesion;
      const T mean = tr / T(3);
      const T dev[3] = {e[0] - m.cohesion - mean, e[1] - m.cohesion - mean, e[2] - m.cohesion - mean};
      const T dn = tsqrt(dev[0] * dev[0] + dev[1] * dev[1] + dev[2] * dev[2]);
      bool changed = false;
      if (tr >= T(0) || dn == T(0)) {   // tension (or on the axis): project to the tip
        if (tr >= T(0)) { e[0] = e[1] = e[2] = m.cohesion; changed = true; }
      } else {
        const T dg = dn + (T(3) * m.lambda + T(2) * m.mu) / (T(2) * m.mu) * tr * m.alpha;
        if (dg > T(0)) {
          const T f = dg / dn;
#pragma unroll
          for (int i = 0; i < 3; ++i) e[i] -= f * dev[i];
          changed = true;
        }
      }
      if (changed || s[2] <= floor_s) {
        T se[3] = {texp(e[0]), texp(e[1]), texp(e[2])};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j)
            Fn[i * 3 + j] = U[i][0] * se[0] * V[j][0] + U[i][1] * se[1] * V[j][1] + U[i][2] * se[2] * V[j][2];
      }
    }
    const T tre = e[0] + e[1] + e[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = T(2) * m.mu * e[i] + m.lambda * tre;
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) F[i] = Fn[i];
  tau[0] = U[0][0] * t[0] * U[0][0] + U[0][1] * t[1] * U[0][1] + U[0][2] * t[2] * U[0][2];
  tau[1] = U[1][0] * t[0] * U[1][0] + U[1][1] * t[1] * U[1][1] + U[1][2] * t[2] * U[1][2];
  tau[2] = U[2][0] * t[0] * U[2][0] + U[2][1] * t[1] * U[2][1] + U[2][2] * t[2] * U[2][2];
  tau[3] = U[0][0] * t[0] * U[1][0] + U[0][1] * t[1] * U[1][1] + U[0][2] * t[2] * U[1][2];
  tau[4] = U[0][0] * t[0] * U[2][0] + U[0][1] * t[1] * U[2][1] + U[0][2] * t[2] * U[2][2];
  tau[5] = U[1][0] * t[0] * U[2][0] + U[1][1] * t[1] * U[2][1] + U[1][2] * t[2] * U[2][2];
}

// CalcFemStateAndForce for the volumetric particles: one thread per physical slot (C, vol, taus are planar by slot:
// coalesced); F lives by volumetric particle id (72 contiguous bytes per particle).
#ifndef GMPM_VOL_THREADS
#define GMPM_VOL_THREADS 128
#endif
#ifndef GMPM_VOL_MIN_CTAS
#define GMPM_VOL_MIN_CTAS 8
#endif
constexpr int kVolThreads = GMPM_VOL_THREADS;
template <typename T>
__global__ void __launch_bounds__(GMPM_VOL_THREADS, GMPM_VOL_MIN_CTAS) volumetric_stress_kernel(uint32_t N, uint32_t n_cloth, const int* __restrict__ pid,
                                         const T* __restrict__ C, const T* __restrict__ vol, T* __restrict__ Fv,
                                         const uint8_t* __restrict__ mat_id, MaterialTable<T> mats, T* __restrict__ taus,
                                         T dt) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= N) return;
  const uint32_t p = (uint32_t)pid[s];
  if (p < n_cloth) return;
  const uint32_t q = p - n_cloth;
  const T v0 = vol[s];
  if (!(v0 > T(0))) {  // inactive spare of the migration (csrc/migrate.cuh): no mass, no stress, F stays I
#pragma unroll
    for (int e = 0; e < 6; ++e) taus[(size_t)e * N + s] = T(0);
    return;
  }
  T Cl[9], F[9], tau[6];
#pragma unroll
  for (int e = 0; e < 9; ++e) {
    Cl[e] = C[(size_t)e * N + s];
    F[e] = Fv[(size_t)q * 9 + e];
  }
  volumetric_update<T>(mats.m[mat_id[q]], Cl, dt, F, tau);
#pragma unroll
  for (int e = 0; e < 9; ++e) Fv[(size_t)q * 9 + e] = F[e];
#pragma unroll
  for (int e = 0; e < 6; ++e) taus[(size_t)e * N + s] = v0 * tau[e];
}

}  // namespace gmpm
