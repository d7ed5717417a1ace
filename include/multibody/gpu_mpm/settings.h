// Drop-in replacement header for the reference's multibody/gpu_mpm/settings.h (settings.h:26-126),
// host-only: the constants below are the DEFAULTS of the runtime parameters of libgmpm_b200
// (gmpm_params in gmpm_c.h), kept under the reference's names for callers that read them.
// Header-only C++17, no CUDA headers needed: everything device-side lives behind the C ABI.
#pragma once

#include <cstddef>
#include <cstdint>

#if defined(GMPM_B200_NO_EIGEN)
#define GMPM_B200_HAVE_EIGEN 0
#elif defined(__has_include)
#if __has_include(<Eigen/Dense>)
#include <Eigen/Dense>
#define GMPM_B200_HAVE_EIGEN 1
#elif __has_include(<eigen3/Eigen/Dense>)
#include <eigen3/Eigen/Dense>
#define GMPM_B200_HAVE_EIGEN 1
#else
#define GMPM_B200_HAVE_EIGEN 0
#endif
#else
#define GMPM_B200_HAVE_EIGEN 0
#endif

namespace drake {
namespace multibody {
namespace gmpm {

#if GMPM_B200_HAVE_EIGEN
template <typename T> using Vec3 = Eigen::Matrix<T, 3, 1>;   // settings.h:31
template <typename T> using Mat3 = Eigen::Matrix<T, 3, 3>;
#else
// Minimal stand-in used where Eigen is not installed (this repository's own tests): three packed
// scalars with the few members the MPM interface touches.
template <typename T>
struct Vec3 {
  T v[3];
  Vec3() : v{T(0), T(0), T(0)} {}
  Vec3(T x, T y, T z) : v{x, y, z} {}
  T& operator[](int i) { return v[i]; }
  const T& operator[](int i) const { return v[i]; }
  T* data() { return v; }
  const T* data() const { return v; }
  static Vec3 Zero() { return Vec3(); }
  Vec3& operator/=(T s) { v[0] /= s; v[1] /= s; v[2] /= s; return *this; }
  template <typename U> Vec3<U> cast() const { return Vec3<U>(U(v[0]), U(v[1]), U(v[2])); }  // Eigen's .cast<U>()
};
#endif
static_assert(sizeof(Vec3<float>) == 3 * sizeof(float) && sizeof(Vec3<double>) == 3 * sizeof(double),
              "Vec3<T> must be three packed scalars: vectors of it are handed to the C ABI as T[3n]");

namespace config {
#if defined(GMPM_B200_SCALAR_DOUBLE)
using GpuT = double;   // the fp64 instantiation of the same algorithms
#else
using GpuT = float;    // settings.h:37 (the reference's product build)
#endif
constexpr int BLOCK_BITS = 2;             // settings.h:48
constexpr int DOMAIN_BITS = 7;            // settings.h:49 (default of gmpm_params::domain_bits)
constexpr int G_DOMAIN_BITS = DOMAIN_BITS;
constexpr int G_DOMAIN_SIZE = 1 << DOMAIN_BITS;
constexpr int G_BOUNDARY_CONDITION = 3;   // settings.h:56
constexpr int G_BLOCK_BITS = BLOCK_BITS;
constexpr int G_BLOCK_SIZE = 1 << BLOCK_BITS;
constexpr int G_BLOCK_VOLUME = 1 << (3 * BLOCK_BITS);
template <class T> constexpr T G_DX = T(1) / T(1 << DOMAIN_BITS);   // settings.h:57
template <class T> constexpr T G_DX_INV = T(1 << DOMAIN_BITS);
template <class T> constexpr T YOUNGS_MODULUS = T(400000);          // settings.h:72-74
template <class T> constexpr T POISSON_RATIO = T(0.3);              // settings.h:76-78
template <class T> constexpr T DENSITY = T(2000);                   // settings.h:80-82
template <class T> constexpr T SDF_FRICTION = T(0.3);               // settings.h:109-111
constexpr int GRAVITY_AXIS = 2;                                     // settings.h:118
template <class T> constexpr T GRAVITY = T(-9.8);                   // settings.h:120-122
}  // namespace config

}  // namespace gmpm
}  // namespace multibody
}  // namespace drake
