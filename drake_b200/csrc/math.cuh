// math.cuh -- small fixed-size matrix helpers and the cloth constitutive model (device only).
// All matrices are row-major.  The algorithms follow the reference so that results agree to
// rounding (math_tools.cuh:113-150, 456-597; cuda_mpm_kernels.cuh:72-181), written here as fully
// unrolled register code for sm_100a.
#pragma once
#include "common.cuh"

namespace gmpm {

template <typename T> __device__ __forceinline__ T tsqrt(T x);
template <> __device__ __forceinline__ float tsqrt<float>(float x) { return sqrtf(x); }
template <> __device__ __forceinline__ double tsqrt<double>(double x) { return sqrt(x); }
template <typename T> __device__ __forceinline__ T tabs(T x) { return x < T(0) ? -x : x; }
template <typename T> __device__ __forceinline__ T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> __device__ __forceinline__ T tmax(T a, T b) { return a > b ? a : b; }

// C[3x3] = A[3x3] * B[3x3]
template <typename T>
__device__ __forceinline__ void mul33(const T* a, const T* b, T* c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      c[i * 3 + k] = a[i * 3 + 0] * b[0 * 3 + k] + a[i * 3 + 1] * b[1 * 3 + k] + a[i * 3 + 2] * b[2 * 3 + k];
}
// C[3x3] = A[3x3] * B[3x3]^T
template <typename T>
__device__ __forceinline__ void mul33T(const T* a, const T* b, T* c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int k = 0; k < 3; ++k)
      c[i * 3 + k] = a[i * 3 + 0] * b[k * 3 + 0] + a[i * 3 + 1] * b[k * 3 + 1] + a[i * 3 + 2] * b[k * 3 + 2];
}
// y = A[3x3] x
template <typename T>
__device__ __forceinline__ void mul33v(const T* a, const T* x, T* y) {
#pragma unroll
  for (int i = 0; i < 3; ++i) y[i] = a[i * 3 + 0] * x[0] + a[i * 3 + 1] * x[1] + a[i * 3 + 2] * x[2];
}
// y = A[3x3]^T x
template <typename T>
__device__ __forceinline__ void mul33Tv(const T* a, const T* x, T* y) {
#pragma unroll
  for (int i = 0; i < 3; ++i) y[i] = a[0 * 3 + i] * x[0] + a[1 * 3 + i] * x[1] + a[2 * 3 + i] * x[2];
}
template <typename T>
__device__ __forceinline__ T det3(const T* m) {
  return m[0] * (m[4] * m[8] - m[7] * m[5]) - m[3] * (m[1] * m[8] - m[7] * m[2]) +
         m[6] * (m[1] * m[5] - m[4] * m[2]);
}
template <typename T>
__device__ __forceinline__ void inverse3(const T* m, T* mi) {
  const T di = T(1) / det3(m);
  mi[0] = (m[4] * m[8] - m[5] * m[7]) * di;
  mi[3] = (m[5] * m[6] - m[3] * m[8]) * di;
  mi[6] = (m[3] * m[7] - m[4] * m[6]) * di;
  mi[1] = (m[2] * m[7] - m[1] * m[8]) * di;
  mi[4] = (m[0] * m[8] - m[2] * m[6]) * di;
  mi[7] = (m[1] * m[6] - m[0] * m[7]) * di;
  mi[2] = (m[1] * m[5] - m[2] * m[4]) * di;
  mi[5] = (m[2] * m[3] - m[0] * m[5]) * di;
  mi[8] = (m[0] * m[4] - m[1] * m[3]) * di;
}
template <typename T>
__device__ __forceinline__ T det2(const T* m) { return m[0] * m[3] - m[1] * m[2]; }
template <typename T>
__device__ __forceinline__ void inverse2(const T* m, T* mi) {
  const T di = T(1) / det2(m);
  mi[0] = m[3] * di;
  mi[1] = -m[1] * di;
  mi[2] = -m[2] * di;
  mi[3] = m[0] * di;
}

// One Givens rotation that zeroes `b` against `a` (math_tools.cuh:470-483).
template <typename T>
__device__ __forceinline__ void givens(T a, T b, T& c, T& s) {
  c = T(1);
  s = T(0);
  const T r = tsqrt(a * a + b * b);
  if (r > T(0)) {
    const T t = T(1) / r;
    c = a * t;
    s = -b * t;
  }
}
template <typename T, int W>
__device__ __forceinline__ void rot_rows(T* m, int r0, int r1, T c, T s) {
#pragma unroll
  for (int j = 0; j < W; ++j) {
    const T t0 = m[r0 * W + j], t1 = m[r1 * W + j];
    m[r0 * W + j] = c * t0 - s * t1;
    m[r1 * W + j] = s * t0 + c * t1;
  }
}
// Givens QR of a 3 x M matrix (M = 2 or 3): A = Q R, rotation order (1,2),(0,1) for column 0 then
// (1,2) for column 1 -- the order of the reference's givens_QR<3,M> (math_tools.cuh:456-510).
template <typename T, int M>
__device__ __forceinline__ void qr3(const T* A, T* Q, T* R) {
#pragma unroll
  for (int i = 0; i < 3 * M; ++i) R[i] = A[i];
  T Qt[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
  T c, s;
  givens(R[1 * M + 0], R[2 * M + 0], c, s);
  rot_rows<T, M>(R, 1, 2, c, s);
  rot_rows<T, 3>(Qt, 1, 2, c, s);
  givens(R[0 * M + 0], R[1 * M + 0], c, s);
  rot_rows<T, M>(R, 0, 1, c, s);
  rot_rows<T, 3>(Qt, 0, 1, c, s);
  givens(R[1 * M + 1], R[2 * M + 1], c, s);
  rot_rows<T, M>(R, 1, 2, c, s);
  rot_rows<T, 3>(Qt, 1, 2, c, s);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) Q[i * 3 + j] = Qt[j * 3 + i];
}

// 2x2 polar decomposition A = U P (math_tools.cuh:512-549).
template <typename T>
__device__ __forceinline__ void polar2(const T* A, T* U, T* P) {
  U[0] = T(1); U[1] = T(0); U[2] = T(0); U[3] = T(1);
  P[0] = A[0]; P[1] = A[1]; P[2] = A[2]; P[3] = A[3];
  if (A[0] == T(0) && A[1] == T(0) && A[2] == T(0) && A[3] == T(0)) return;
  const T detA = det2(A);
  const T adetA = tabs(detA);
  T B[4] = {A[0] + A[3], A[1] - A[2], A[2] - A[1], A[3] + A[0]};
  if (detA < T(0)) {
    B[0] = A[0] - A[3];
    B[1] = A[1] + A[2];
    B[2] = A[2] + A[1];
    B[3] = A[3] - A[0];
  }
  const T k = T(1) / tsqrt(tabs(det2(B)));
  U[0] = B[0] * k; U[1] = B[1] * k; U[2] = B[2] * k; U[3] = B[3] * k;
  P[0] = (A[0] * A[0] + A[2] * A[2] + adetA) * k;
  P[1] = (A[0] * A[1] + A[2] * A[3]) * k;
  P[2] = P[1];
  P[3] = (A[1] * A[1] + A[3] * A[3] + adetA) * k;
}

// Rotation factor R = U V^T of the 2x2 SVD (math_tools.cuh:551-597 followed by
// cuda_mpm_kernels.cuh:76-78).  Only U V^T is needed by the fixed-corotated stress, and
// (R_polar V) V^T is evaluated the same way the reference does (two products) to keep rounding.
template <typename T>
__device__ __forceinline__ void svd2_rotation(const T* A, T* Rot) {
  T Rp[4], S[4];
  polar2(A, Rp, S);
  T c, s, s1, s2;
  if (tabs(S[1]) < T(1e-5)) {
    c = T(1); s = T(0); s1 = S[0]; s2 = S[3];
  } else {
    const T tao = T(.5) * (S[0] - S[3]);
    const T w = tsqrt(tao * tao + S[1] * S[1]);
    const T t = tao > T(0) ? S[1] / (tao + w) : S[1] / (tao - w);
    c = T(1) / tsqrt(t * t + T(1));
    s = -t * c;
    s1 = c * c * S[0] - T(2) * c * s * S[1] + s * s * S[3];
    s2 = s * s * S[0] + T(2) * c * s * S[1] + c * c * S[3];
  }
  T V[4];
  if (s1 < s2) {
    V[0] = -s; V[1] = c; V[2] = -c; V[3] = -s;
  } else {
    V[0] = c; V[1] = s; V[2] = -s; V[3] = c;
  }
  T U[4];
  U[0] = Rp[0] * V[0] + Rp[1] * V[2];
  U[1] = Rp[0] * V[1] + Rp[1] * V[3];
  U[2] = Rp[2] * V[0] + Rp[3] * V[2];
  U[3] = Rp[2] * V[1] + Rp[3] * V[3];
  Rot[0] = U[0] * V[0] + U[1] * V[1];
  Rot[1] = U[0] * V[2] + U[1] * V[3];
  Rot[2] = U[2] * V[0] + U[3] * V[1];
  Rot[3] = U[2] * V[2] + U[3] * V[3];
}

// 2-D fixed-corotated first Piola stress (cuda_mpm_kernels.cuh:72-86).
template <typename T>
__device__ __forceinline__ void fixed_corotated_pk1_2d(const T* F, T mu, T lambda, T* P) {
  T R[4], Fi[4];
  svd2_rotation(F, R);
  const T J = det2(F);
  inverse2(F, Fi);
  const T lj = lambda * (J - T(1)) * J;
  P[0] = T(2) * mu * (F[0] - R[0]) + lj * Fi[0];
  P[1] = T(2) * mu * (F[1] - R[1]) + lj * Fi[2];
  P[2] = T(2) * mu * (F[2] - R[2]) + lj * Fi[1];
  P[3] = T(2) * mu * (F[3] - R[3]) + lj * Fi[3];
}

// Cloth energy derivative dphi/dF (cuda_mpm_kernels.cuh:88-144): in-plane fixed-corotated on the
// R factor of F = QR plus the shear (gamma) and normal (k) terms.
template <typename T>
__device__ __forceinline__ void cloth_pk1(const Consts<T>& k, const T* F, T* P) {
  T Q[9], R[9];
  qr3<T, 3>(F, Q, R);
  const T Rh[4] = {R[0], R[1], R[3], R[4]};
  T p2[4];
  fixed_corotated_pk1_2d(Rh, k.mu, k.lambda, p2);
  const T Ph[9] = {p2[0], p2[1], T(0), p2[2], p2[3], T(0), T(0), T(0), T(0)};
  T Pplane[9];
  mul33(Q, Ph, Pplane);
  const T gp = k.gamma;
  T fp = T(0);
  if (R[8] < T(1)) fp = -k.k_normal * (T(1) - R[8]) * (T(1) - R[8]);
  T A[9];
  A[0] = gp * R[2] * R[2];
  A[1] = gp * R[2] * R[5];
  A[2] = gp * R[8] * R[2];
  A[4] = gp * R[5] * R[5];
  A[5] = gp * R[8] * R[8];
  A[8] = fp * R[8];
  A[3] = A[1];
  A[6] = A[2];
  A[7] = A[5];
  T Ri[9], QA[9], Pn[9];
  inverse3(R, Ri);
  mul33(Q, A, QA);
  mul33T(QA, Ri, Pn);
#pragma unroll
  for (int i = 0; i < 9; ++i) P[i] = Pplane[i] + Pn[i];
}

// Return mapping on F = QR (cuda_mpm_kernels.cuh:146-181).
template <typename T>
__device__ __forceinline__ void cloth_project_strain(const Consts<T>& k, T* F) {
  T Q[9], R[9];
  qr3<T, 3>(F, Q, R);
  if (k.gamma == T(0)) {
    R[8] = tmin(R[8], T(1));
    R[2] = T(0);
    R[5] = T(0);
  } else if (R[8] > T(1)) {
    R[8] = tmin(R[8], T(1));
    R[2] = T(0);
    R[5] = T(0);
  } else if (R[8] <= T(0)) {
    R[2] = T(0);
    R[5] = T(0);
    R[8] = tmax(R[8], T(-1));
  } else {
    const T rr = R[2] * R[2] + R[5] * R[5];
    const T gk = k.gamma / k.k_normal;
    const T zz = k.c_f * (R[8] - T(1)) * (R[8] - T(1));
    const T f = (gk * gk) * rr - (zz * zz);
    if (f > T(0)) {
      const T c = zz / (gk * tsqrt(rr));
      R[2] *= c;
      R[5] *= c;
    }
  }
  mul33(Q, R, F);
}

// Orthonormal frame with row `axis` = u (math_tools.cuh:599-638, restating Drake's
// RotationMatrix::MakeFromOneUnitVector).
template <typename T>
__device__ __forceinline__ void frame_from_unit_vector(const T* u, int axis, T* J) {
  int i = 0;
  if (tabs(u[1]) < tabs(u[i])) i = 1;
  if (tabs(u[2]) < tabs(u[i])) i = 2;
  const int j = (i + 1) % 3, k = (j + 1) % 3;
  const T mag = tsqrt(T(1) - u[i] * u[i]);
  const T r = T(1) / mag;
  const T s = -r * u[i];
  J[axis * 3 + 0] = u[0];
  J[axis * 3 + 1] = u[1];
  J[axis * 3 + 2] = u[2];
  const int vi = (axis + 1) % 3, wi = (axis + 2) % 3;
  J[vi * 3 + i] = T(0);
  J[vi * 3 + j] = -r * u[k];
  J[vi * 3 + k] = r * u[j];
  J[wi * 3 + i] = mag;
  J[wi * 3 + j] = s * u[j];
  J[wi * 3 + k] = s * u[k];
}

}  // namespace gmpm
