// ref_host_probe.cu -- runs the reference's OWN __host__ __device__ helpers on the host CPU (they
// compile for host under nvcc) on seeded inputs and prints inputs + outputs as JSON with exact
// hex-float encoding.  tests/golden/make_ref_helper_golden.py runs it in the build container (where
// /root/reference exists) and commits the result as tests/golden/ref_helpers.json; the CPU oracle
// is then pinned against it bit-for-bit.  TEST INFRASTRUCTURE ONLY.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cassert>

#include "multibody/gpu_mpm/cuda_mpm_kernels.cuh"

using namespace drake::multibody::gmpm;

static uint64_t g_seed = 20261104ull;
static double urand() {  // splitmix64 -> [0,1)
  uint64_t z = (g_seed += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

template <typename T>
static void arr(const char* name, const T* v, int n, bool last = false) {
  printf("\"%s\": [", name);
  for (int i = 0; i < n; ++i) printf("\"%a\"%s", (double)v[i], i + 1 < n ? ", " : "");
  printf("]%s", last ? "" : ", ");
}

template <typename T>
static void fp_cases(const char* tag, int n_cases) {
  printf("\"%s\": [\n", tag);
  for (int c = 0; c < n_cases; ++c) {
    // deformation gradients near identity / rotated / compressed / inverted-normal
    T F[9];
    const double amp = (c % 4 == 0) ? 0.05 : (c % 4 == 1) ? 0.3 : (c % 4 == 2) ? 1.0 : 2.0;
    for (int i = 0; i < 9; ++i) F[i] = T((i % 4 == 0 ? 1.0 : 0.0) + amp * (urand() - 0.5));
    if (c % 7 == 3) F[8] = T(-0.2);             // inversion branch input
    if (c % 5 == 1) { F[1] = F[3]; }            // near-symmetric in-plane -> small S[1]
    T Q[9], R[9];
    givens_QR<3, 3, T>(F, Q, R);
    T A32[6] = {F[0], F[1], F[3], F[4], F[6], F[7]}, Q32[9], R32[6];
    givens_QR<3, 2, T>(A32, Q32, R32);
    T A2[4] = {F[0], F[1], F[3], F[4]}, U2[4], S2[4], V2[4], PU[4], PP[4];
    svd2x2(A2, U2, S2, V2);
    polar_decompose2x2(A2, PU, PP);
    T pk2[4];
    fixed_corotated_PK1_2D(A2, pk2);
    T P[9];
    compute_dphi_dF(F, P);
    T Fp[9];
    for (int i = 0; i < 9; ++i) Fp[i] = F[i];
    project_strain(Fp);
    T Finv[9];
    inverse3(F, Finv);
    // unit vector for make_from_one_unit_vector
    T u[3] = {T(urand() - 0.5), T(urand() - 0.5), T(urand() - 0.5)};
    if (c % 6 == 0) { u[0] = 0; u[1] = 0; u[2] = T(c % 12 == 0 ? 1 : -1); }
    normalize<3, T>(u);
    T J[9];
    make_from_one_unit_vector(u, 2, J);
    printf("  {");
    arr("F", F, 9); arr("Q", Q, 9); arr("R", R, 9); arr("Q32", Q32, 9); arr("R32", R32, 6);
    arr("U2", U2, 4); arr("S2", S2, 4); arr("V2", V2, 4); arr("polarU", PU, 4); arr("polarP", PP, 4);
    arr("pk2", pk2, 4); arr("P", P, 9); arr("Fproj", Fp, 9); arr("Finv", Finv, 9);
    arr("u", u, 3); arr("J", J, 9, true);
    printf("}%s\n", c + 1 < n_cases ? "," : "");
  }
  printf("]");
}

int main() {
  printf("{\n\"source\": \"reference __host__ __device__ helpers run on host via nvcc (cuda_mpm_kernels.cuh, math_tools.cuh)\",\n");
  // integer known answers
  printf("\"cell_index\": [\n");
  const int n_int = 64;
  for (int c = 0; c < n_int; ++c) {
    uint32_t x = (uint32_t)(urand() * 128), y = (uint32_t)(urand() * 128), z = (uint32_t)(urand() * 128);
    if (c == 0) { x = 1; y = 2; z = 3; }
    if (c == 1) { x = 5; y = 9; z = 127; }
    if (c == 2) { x = 64; y = 64; z = 96; }
    if (c == 3) { x = 127; y = 127; z = 127; }
    if (c == 4) { x = 0; y = 0; z = 0; }
    if (c >= 48) { x = (uint32_t)(urand() * 512); y = (uint32_t)(urand() * 512); z = (uint32_t)(urand() * 512); }
    uint32_t key = cell_index(x, y, z);
    uint3 inv = inverse_cell_index(key);
    uint32_t mort = morton_code({x >> 2, y >> 2, z >> 2});
    printf("  {\"xyz\": [%u, %u, %u], \"key\": %u, \"inv\": [%u, %u, %u], \"block_morton\": %u, \"color\": %u}%s\n",
           x, y, z, key, inv.x, inv.y, inv.z, mort, get_color_mask<uint32_t>(x, y, z), c + 1 < n_int ? "," : "");
  }
  printf("],\n\"morton_31_31_31\": %u,\n", morton_code({31u, 31u, 31u}));
  fp_cases<float>("f32", 48);
  printf(",\n");
  fp_cases<double>("f64", 48);
  printf("\n}\n");
  return 0;
}
