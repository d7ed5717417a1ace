/*
 * gmpm_oracle.h -- CPU restatement of g1n0st/drake's multibody/gpu_mpm hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under drake_b200/ may include, link or call
 * this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker / reported CPU baseline.
 *
 * The reference ships this path ONLY as CUDA kernels (no CPU implementation, no
 * golden vectors: multibody/gpu_mpm/cuda_mpm_test.cc:61 asserts a non-null
 * pointer and nothing else).  This oracle restates every kernel body as a plain
 * serial loop in particle / face / cell index order, with the same expression
 * forms (incl. the reference's accidental double promotions such as `1. / x`).
 * Each function cites the reference file:line it follows.
 *
 * Pinning: (1) integer known answers probed from the reference's own
 * __host__ __device__ helpers (cell_index / morton_code values);
 * (2) fp known answers of compute_dphi_dF / project_strain probed on host from
 * the reference helpers; (3) the contact formulas against the
 * double-precision CPU SAP Hunt-Crossley formulas in
 * multibody/contact_solvers/sap/sap_hunt_crossley_constraint.cc:67-267;
 * (4) on the GPU box, against the reference's own kernels compiled for sm_100
 * from /root/reference (oracle/_ref, see oracle/Makefile).
 *
 * Two scalar instantiations: suffix _f32 (the reference's product build,
 * config::GpuT = float, settings.h:37) and _f64 (the same templates at double).
 * Build with -ffp-contract=off so the arithmetic is the literal expression tree.
 */
#ifndef GMPM_ORACLE_H_
#define GMPM_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Runtime copy of the compile-time constants in settings.h:37-126.  Defaults
 * (ogmpm_default_params) are the reference's values. */
typedef struct ogmpm_params {
  int domain_bits;      /* settings.h:49 DOMAIN_BITS = 7 -> 128^3 cells           */
  int boundary_cells;   /* settings.h:56 G_BOUNDARY_CONDITION = 3                 */
  int gravity_axis;     /* settings.h:118 GRAVITY_AXIS = 2                        */
  int reserved;
  double youngs;        /* settings.h:72-74   4e5                                 */
  double poisson;       /* settings.h:76-78   0.3                                 */
  double density;       /* settings.h:80-82   2000                                */
  double gamma;         /* settings.h:84-86   0                                   */
  double k_normal;      /* settings.h:91-93   K = 1e5                             */
  double v_blend;       /* settings.h:98-100  V = 0.8                             */
  double c_f;           /* settings.h:102-104 c_F = 0                             */
  double sdf_friction;  /* settings.h:109-111 0.3                                 */
  double gravity;       /* settings.h:120-122 -9.8                                */
  double epsv;          /* settings.h:124-126 1e-3                                */
  double dx;            /* settings.h:57 G_DX; 0 = 1/2^domain_bits (the reference) */
} ogmpm_params;

void ogmpm_default_params(ogmpm_params* p);

/* integer helpers, cuda_mpm_kernels.cuh:296-363 */
uint32_t ogmpm_expand_bits(uint32_t v);
uint32_t ogmpm_contract_bits(uint32_t v);
uint32_t ogmpm_morton_code(uint32_t x, uint32_t y, uint32_t z);
uint32_t ogmpm_cell_index(uint32_t xi, uint32_t yi, uint32_t zi);
void ogmpm_inverse_cell_index(uint32_t index, uint32_t* xyz);
/* radix_sort.cuh:202-286 semantics: stable LSD sort of (key,id) on bits [0,num_bit) */
void ogmpm_stable_sort_low_bits(const uint32_t* keys_in, const uint32_t* ids_in, size_t n,
                                int num_bit, uint32_t* keys_out, uint32_t* ids_out);

typedef struct ogmpm_contact_stats {
  int iterations;            /* `count`, cuda_mpm_solver.cu:241,570   */
  int grid_dofs;             /* `grid_DoFs`, :251,564                 */
  int line_search_evals;     /* sum of s_line_search_cnts, :530       */
  double residual;           /* final `norm_dir`, :569                */
  double energy;             /* last `global_E1`, :531                */
} ogmpm_contact_stats;

#define OGMPM_DECLARE(SUF, T)                                                                   \
  typedef struct ogmpm_state_##SUF ogmpm_state_##SUF;                                           \
  ogmpm_state_##SUF* ogmpm_create_##SUF(const ogmpm_params* p);                                 \
  void ogmpm_destroy_##SUF(ogmpm_state_##SUF* s);                                               \
  /* cuda_mpm_model.cu:16-33 */                                                                 \
  void ogmpm_add_qr_cloth_##SUF(ogmpm_state_##SUF* s, const T* pos, const T* vel, size_t nv,    \
                                const int* tri, size_t nf);                                     \
  /* extension (gmpm_oracle_vol.inc; parity unpinned): particle-only bodies with a volumetric    */\
  /* material; model 1 fixed corotated, 2 StVK-Hencky, 3 Drucker-Prager                          */\
  void ogmpm_add_particles_##SUF(ogmpm_state_##SUF* s, const T* pos, const T* vel, const T* vol, \
                                 size_t n, int model, double youngs, double poisson,             \
                                 double friction_angle_deg, double cohesion);                    \
  size_t ogmpm_n_volumetric_##SUF(const ogmpm_state_##SUF* s);                                   \
  size_t ogmpm_migrate_out_##SUF(ogmpm_state_##SUF* s, int axis, T lo, T hi, const T* park,      \
                                 T* records, int* pids, size_t cap);                             \
  int ogmpm_migrate_in_##SUF(ogmpm_state_##SUF* s, const T* records, size_t n, int* pids);       \
  void ogmpm_svd3_##SUF(const T* F, T* U9, T* sig, T* V9);                                       \
  void ogmpm_volumetric_update_##SUF(int model, double youngs, double poisson,                   \
                                     double friction_angle_deg, double cohesion, const T* C,     \
                                     T dt, T* F, T* tau9);                                       \
  /* cuda_mpm_model.cu:36-122 + cuda_mpm_kernels.cuh:13-70 */                                   \
  void ogmpm_finalize_##SUF(ogmpm_state_##SUF* s);                                              \
  /* cuda_mpm_solver.cu:18-70 */                                                                \
  void ogmpm_rebuild_mapping_##SUF(ogmpm_state_##SUF* s, int sort);                             \
  /* cuda_mpm_solver.cu:73-84 + cuda_mpm_kernels.cuh:183-294 */                                 \
  void ogmpm_calc_fem_##SUF(ogmpm_state_##SUF* s, T dt);                                        \
  /* cuda_mpm_solver.cu:87-105 + cuda_mpm_kernels.cuh:418-543,585-602 */                        \
  void ogmpm_p2g_##SUF(ogmpm_state_##SUF* s, T dt);                                             \
  /* cuda_mpm_solver.cu:108-151 + cuda_mpm_kernels.cuh:545-583,632-796 */                       \
  void ogmpm_update_grid_##SUF(ogmpm_state_##SUF* s, int mpm_bc);                               \
  /* cuda_mpm_solver.cu:154-161 + cuda_mpm_kernels.cuh:798-924 */                               \
  void ogmpm_g2p_##SUF(ogmpm_state_##SUF* s, T dt);                                             \
  /* cuda_mpm_solver.cu:194-212 + cuda_mpm_kernels.cuh:926-938 */                               \
  void ogmpm_set_contacts_##SUF(ogmpm_state_##SUF* s, size_t n, const uint32_t* mpm_id,         \
                                const uint32_t* rigid_id, const T* dist, const T* normal,       \
                                const T* pos, const T* rigid_v, const T* rigid_p);              \
  /* cuda_mpm_solver.cu:215-621 */                                                              \
  void ogmpm_update_contact_##SUF(ogmpm_state_##SUF* s, T dt, T mu, T stiffness, T damping,     \
                                  int exact_line_search, int max_iterations,                    \
                                  ogmpm_contact_stats* stats);                                  \
  /* cuda_mpm_model.cu:320-345 */                                                               \
  void ogmpm_set_num_bodies_##SUF(ogmpm_state_##SUF* s, size_t n);                              \
  void ogmpm_body_impulses_##SUF(const ogmpm_state_##SUF* s, T* tau, T* f);                     \
  /* cuda_mpm_model.cu:245-265 */                                                               \
  void ogmpm_dump_cpu_state_##SUF(const ogmpm_state_##SUF* s, T* vert_pos, int* tri);           \
  /* per-contact local terms, cuda_mpm_kernels.cuh:956-1040 and the cost l(v) :1425-1435 */    \
  void ogmpm_contact_grad_hess_##SUF(const ogmpm_params* p, T phi0, T dt, T stiffness,          \
                                     T damping, T mu, const T* v0, const T* v_next, T* hess,    \
                                     T* grad);                                                  \
  T ogmpm_contact_cost_##SUF(const ogmpm_params* p, T phi0, T dt, T stiffness, T damping, T mu, \
                             const T* v0, const T* v);                                          \
  /* math_tools.cuh:599-638 */                                                                  \
  void ogmpm_make_from_one_unit_vector_##SUF(const T* u, int axis, T* J);                       \
  /* cuda_mpm_kernels.cuh:88-181 */                                                             \
  void ogmpm_compute_dphi_dF_##SUF(const ogmpm_params* p, const T* F, T* P);                    \
  void ogmpm_project_strain_##SUF(const ogmpm_params* p, T* F);                                 \
  /* math_tools.cuh:456-597 */                                                                  \
  void ogmpm_givens_qr33_##SUF(const T* A, T* Q, T* R);                                         \
  void ogmpm_svd2x2_##SUF(const T* A, T* U, T* sig, T* V);                                      \
  /* sizes + raw views (pointers into the oracle state; valid until next call) */              \
  size_t ogmpm_n_verts_##SUF(const ogmpm_state_##SUF* s);                                       \
  size_t ogmpm_n_faces_##SUF(const ogmpm_state_##SUF* s);                                       \
  size_t ogmpm_n_particles_##SUF(const ogmpm_state_##SUF* s);                                   \
  size_t ogmpm_n_cells_##SUF(const ogmpm_state_##SUF* s);                                       \
  size_t ogmpm_n_contacts_##SUF(const ogmpm_state_##SUF* s);                                    \
  uint32_t ogmpm_touched_count_##SUF(const ogmpm_state_##SUF* s);                               \
  const void* ogmpm_view_##SUF(const ogmpm_state_##SUF* s, const char* name);                   \
  /* OpenMP variant of one no-contact substep (keys, fem, p2g, grid, g2p) used only as the */  \
  /* timed CPU baseline; same arithmetic, atomics for the scatters. */                          \
  void ogmpm_substep_omp_##SUF(ogmpm_state_##SUF* s, T dt, int mpm_bc, int threads);

OGMPM_DECLARE(f32, float)
OGMPM_DECLARE(f64, double)

#ifdef __cplusplus
}
#endif
#endif  /* GMPM_ORACLE_H_ */
