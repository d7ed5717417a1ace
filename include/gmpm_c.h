/*
 * gmpm_c.h -- C ABI of libgmpm_b200.so: the B200-native (sm_100a) MPM per-timestep hot path.
 *
 * The reference (g1n0st/drake, multibody/gpu_mpm) has no FFI of its own: the path is reached through
 * two C++ class templates, GpuMpmState<T> (cuda_mpm_model.cuh:37-260) and GpuMpmSolver<T>
 * (cuda_mpm_solver.cuh:19-35), called from DeformableDriver::CalcAbstractStates
 * (multibody/plant/deformable_driver.h:221-271).  Every entry point below replaces exactly one
 * method of those two classes (cited per function); include/multibody/gpu_mpm/ holds header-only
 * C++ shims that re-create the two classes on top of this ABI so deformable_driver.h compiles
 * unchanged.
 *
 * Conventions
 *   - plain pointers and sizes only; every buffer argument is a HOST pointer borrowed for the call
 *     and laid out as the reference lays it out (interleaved xyz, row-major 3x3, visible order);
 *   - one suffix per scalar: _f32 (the reference's product build, settings.h:37) and _f64;
 *   - every function returns 0 on success and a negative gmpm_status otherwise; the message is
 *     available from gmpm_last_error() (thread-local).  The reference has no error reporting
 *     (CUDA_SAFE_CALL is a no-op unless DEBUG=1, settings.h:11-25);
 *   - there is NO CPU fallback: without a CUDA device gmpm_create_* fails with GMPM_ERR_NO_DEVICE;
 *   - all work of one state is issued on one private CUDA stream; calls are asynchronous unless
 *     they return data to the host.  Not re-entrant per state (same as the reference).
 */
#ifndef GMPM_C_H_
#define GMPM_C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gmpm_status {
  GMPM_OK = 0,
  GMPM_ERR_INVALID = -1,   /* bad argument / call order                    */
  GMPM_ERR_NO_DEVICE = -2, /* no usable CUDA device (no CPU fallback)      */
  GMPM_ERR_CUDA = -3,      /* a CUDA runtime call or kernel failed         */
  GMPM_ERR_OOM = -4,       /* device allocation failed                     */
  GMPM_ERR_UNKNOWN_NAME = -5
} gmpm_status;

/* Runtime copy of the reference's compile-time constants (settings.h:37-126).  gmpm_default_params
 * fills in the reference's values.  domain_bits may be raised (8, 9) for the larger scenes. */
typedef struct gmpm_params {
  int domain_bits;     /* settings.h:49  DOMAIN_BITS = 7 -> 128^3 cells, 4^3-cell blocks            */
  int boundary_cells;  /* settings.h:56  G_BOUNDARY_CONDITION = 3                                   */
  int gravity_axis;    /* settings.h:118 GRAVITY_AXIS = 2                                           */
  int device;          /* CUDA device ordinal; -1 = current device                                  */
  double youngs;       /* settings.h:72-74   4e5                                                    */
  double poisson;      /* settings.h:76-78   0.3                                                    */
  double density;      /* settings.h:80-82   2000                                                   */
  double gamma;        /* settings.h:84-86   0                                                      */
  double k_normal;     /* settings.h:91-93   1e5                                                    */
  double v_blend;      /* settings.h:98-100  0.8                                                    */
  double c_f;          /* settings.h:102-104 0                                                      */
  double sdf_friction; /* settings.h:109-111 0.3                                                    */
  double gravity;      /* settings.h:120-122 -9.8                                                   */
  double epsv;         /* settings.h:124-126 1e-3                                                   */
  double dx;           /* settings.h:57 G_DX.  0 = 1 / 2^domain_bits (unit cube, the reference).  A positive */
                       /* value decouples the cell size from the cell count: the domain is the cube        */
                       /* [0, dx * 2^domain_bits]^3 (multi-GPU scenes keep dx = 1/128 on a 256^3 grid)     */
} gmpm_params;

/* Extension (absent upstream): materials of particle-only (volumetric) bodies.  The reference has
 * one law, the QR cloth with compile-time constants (cuda_mpm_kernels.cuh:72-181, settings.h:72-104). */
enum gmpm_material_model {
  GMPM_MAT_FIXED_COROTATED = 1, /* Stomakhin et al. 2013                                           */
  GMPM_MAT_STVK_HENCKY = 2,     /* St.Venant-Kirchhoff with Hencky strain (Klar et al. 2016)       */
  GMPM_MAT_DRUCKER_PRAGER = 3   /* StVK-Hencky + Drucker-Prager return mapping (Klar et al. 2016)  */
};
typedef struct gmpm_material {
  int model;                 /* gmpm_material_model */
  int reserved;
  double youngs, poisson;
  double friction_angle_deg; /* Drucker-Prager */
  double cohesion;           /* Drucker-Prager: log-strain shift of the cone tip (0 = dry sand) */
} gmpm_material;

/* Result of one UpdateContact call (the reference prints these, cuda_mpm_solver.cu:577-582). */
typedef struct gmpm_contact_stats {
  int iterations;        /* Newton iterations (`count`, cuda_mpm_solver.cu:241,570)            */
  int grid_dofs;         /* active grid nodes (`grid_DoFs`, :251,564)                          */
  int line_search_evals; /* total line-search evaluations (:530)                               */
  int reserved;
  double residual;       /* final sqrt(sum |Dir|^2)/DoFs (:569)                                */
  double energy;         /* last accepted E1 (:531)                                            */
} gmpm_contact_stats;

/* Result of gmpm_run_driver_substeps. */
typedef struct gmpm_driver_stats {
  int substeps;
  int contact_queries;       /* substeps in which the contact pairs were regenerated */
  int newton_iterations;     /* summed over the substeps */
  int newton_iterations_max; /* largest count of one substep */
  uint64_t contact_pairs;    /* summed over the substeps */
  double last_residual;
} gmpm_driver_stats;

/* Analytic rigid geometry for device-side contact-pair generation (gmpm_generate_contact_pairs). */
enum { GMPM_GEOM_BOX = 0, GMPM_GEOM_SPHERE = 1 };
typedef struct gmpm_analytic_geometry {
  int kind;            /* GMPM_GEOM_BOX: axis-aligned [lo, hi];  GMPM_GEOM_SPHERE: centre lo, radius hi[0]   */
  uint32_t rigid_id;   /* MpmParticleContactPairs::non_mpm_id (cpu_mpm_model.h:77)                        */
  double lo[3], hi[3];
  double velocity[3];  /* rigid_v: velocity of the body at the contact points (translating body)          */
  double p_WB[3];      /* rigid_p_WB: body origin, the lever arm reference of the contact torque           */
} gmpm_analytic_geometry;

void gmpm_default_params(gmpm_params* p);
const char* gmpm_last_error(void);
/* Library build string, e.g. "gmpm_b200 sm_100a". */
const char* gmpm_version(void);
/* Number of CUDA devices visible (0 -> every create fails). */
int gmpm_device_count(void);
/* GpuMpmSolver<T>::GpuSync, cuda_mpm_solver.cu:164-166: device-wide barrier (all states' streams). */
int gmpm_device_sync(void);

#define GMPM_DECLARE(SUF, T)                                                                        \
  typedef struct gmpm_state_##SUF gmpm_state_##SUF;                                                 \
  /* GpuMpmState<T>() */                                                                            \
  int gmpm_create_##SUF(const gmpm_params* p, gmpm_state_##SUF** out);                              \
  /* GpuMpmState<T>::Destroy, cuda_mpm_model.cu:125-242 */                                          \
  int gmpm_destroy_##SUF(gmpm_state_##SUF* s);                                                      \
  /* GpuMpmState<T>::AddQRCloth, cuda_mpm_model.cu:16-33.  tri indexes this call's vertices. */    \
  int gmpm_add_qr_cloth_##SUF(gmpm_state_##SUF* s, const T* pos, const T* vel, size_t n_verts,      \
                              const int* tri, size_t n_tris);                                       \
  /* Extension: n particles that carry their own deformation gradient (F = I) and a volumetric       */\
  /* material; pos, vel 3n, rest volume n.  Mass = volume * params.density.  Generalises the          */\
  /* AddQRCloth entry point (cuda_mpm_model.cu:16-33); their pids follow the cloth's [faces | verts]. */\
  int gmpm_add_particles_##SUF(gmpm_state_##SUF* s, const T* pos, const T* vel, const T* volume,    \
                               size_t n, const gmpm_material* material);                            \
  /* Extension (multi-GPU): migration of volumetric particles between the states of a      */\
  /* position-owned decomposition.  A state is given spare INACTIVE particles (gmpm_add_particles      */\
  /* with volume 0).  migrate_out copies out the active volumetric particles whose coordinate `axis`   */\
  /* lies outside [lo, hi) -- 26 scalars each: pos 3, vel 3, C 9, volume, F 9, material index -- with  */\
  /* their local particle ids, sorted by id, and turns their slots inactive (parked at a scattered    */\
  /* point of the box park[6] = lo xyz, hi xyz: choose it inside the state's own slab);                */\
  /* at most cap (the others stay for the next call).  migrate_in hands n such records to inactive     */\
  /* slots (pids <- the local ids they took; may be null).  Both belong between GridToParticle and     */\
  /* the next RebuildMapping; all states of a scene must declare their materials in the same order.    */\
  int gmpm_migrate_out_##SUF(gmpm_state_##SUF* s, int axis, T lo, T hi, const T* park, T* records,  \
                             int* pids, size_t cap, size_t* n);                                     \
  int gmpm_migrate_in_##SUF(gmpm_state_##SUF* s, const T* records, size_t n, int* pids);            \
  /* Dry run of migrate_out, so that a migration can be refused BEFORE anything is taken out of a     */\
  /* state (migrate_out destroys the leavers at the source; if their destination then has no spare   */\
  /* slot left they are lost).  Changes nothing.  leavers[0 .. n_edges] <- the number of active       */\
  /* volumetric particles outside [lo, hi) -- NOT limited by any cap, unlike what one migrate_out     */\
  /* call returns -- split by destination slab: destination d = the number of edges <= the particle's */\
  /* coordinate (edges[n_edges] ascending, doubles, <= 63).  *n_free <- inactive slots of this state.     */\
  /* The host sums the leavers per destination over all ranks and compares with n_free there.         */\
  int gmpm_migrate_count_##SUF(gmpm_state_##SUF* s, int axis, T lo, T hi, const double* edges,       \
                               size_t n_edges, size_t* leavers, size_t* n_free);                      \
  /* GpuMpmState<T>::Finalize, cuda_mpm_model.cu:36-122 (+ initialize_fem_state_kernel) */          \
  int gmpm_finalize_##SUF(gmpm_state_##SUF* s);                                                     \
  /* GpuMpmSolver<T>::RebuildMapping, cuda_mpm_solver.cu:18-70 */                                   \
  int gmpm_rebuild_mapping_##SUF(gmpm_state_##SUF* s, int sort);                                    \
  /* GpuMpmSolver<T>::CalcFemStateAndForce, cuda_mpm_solver.cu:73-84 */                             \
  int gmpm_calc_fem_##SUF(gmpm_state_##SUF* s, T dt);                                               \
  /* GpuMpmSolver<T>::ParticleToGrid, cuda_mpm_solver.cu:87-105 */                                  \
  int gmpm_p2g_##SUF(gmpm_state_##SUF* s, T dt);                                                    \
  /* GpuMpmSolver<T>::UpdateGrid, cuda_mpm_solver.cu:108-151 */                                     \
  int gmpm_update_grid_##SUF(gmpm_state_##SUF* s, int mpm_bc);                                      \
  /* GpuMpmSolver<T>::GridToParticle, cuda_mpm_solver.cu:154-161 */                                 \
  int gmpm_g2p_##SUF(gmpm_state_##SUF* s, T dt);                                                    \
  /* GpuMpmSolver<T>::GpuSync, cuda_mpm_solver.cu:164-166 */                                        \
  int gmpm_sync_##SUF(gmpm_state_##SUF* s);                                                         \
  /* GpuMpmSolver<T>::SyncParticleStateToCpu, cuda_mpm_solver.cu:187-191: positions_host(), 3N */  \
  int gmpm_sync_particle_state_to_cpu_##SUF(gmpm_state_##SUF* s, T* positions_host);                \
  /* Same snapshot, not blocking: positions as of THIS call are gathered in stream order and copied */\
  /* on a copy stream into a library-owned PINNED mirror while later solver calls run (the driver   */\
  /* reads positions_host() only when it builds the contact pairs, deformable_driver.h:243 vs :247).*/\
  int gmpm_sync_particle_state_to_cpu_async_##SUF(gmpm_state_##SUF* s);                             \
  /* Extension: declares the axis-aligned bounds of the rigid geometry the host   */\
  /* contact query (deformable_driver.h:120-194) tests against: n boxes, 6 scalars each (lo xyz, hi   */\
  /* xyz), inflated by the caller as it sees fit.  While bounds are set the ASYNC snapshot moves only  */\
  /* the particles inside one of them over PCIe; positions_host() of every other particle reads 1e30  */\
  /* (any signed-distance query is positive there), so the host loop builds the same pairs as from    */\
  /* the full snapshot.  n = 0: only a count crosses.  *d2h_bytes (may be null) of                     */\
  /* gmpm_positions_host_info <- bytes the last snapshot moved.  clear restores the full copy.        */\
  int gmpm_set_query_bounds_##SUF(gmpm_state_##SUF* s, const T* lo_hi, size_t n_boxes);             \
  int gmpm_clear_query_bounds_##SUF(gmpm_state_##SUF* s);                                           \
  int gmpm_snapshot_d2h_bytes_##SUF(gmpm_state_##SUF* s, uint64_t* bytes);                          \
  /* positions_host(): waits for the last async snapshot; *out = pinned mirror, 3N scalars, visible */\
  /* order, valid until the next snapshot or destroy (NULL before the first snapshot).              */\
  int gmpm_positions_host_##SUF(gmpm_state_##SUF* s, const T** out);                                \
  /* GpuMpmSolver<T>::CopyContactPairs, cuda_mpm_solver.cu:194-212 (+ ReallocateContacts) */        \
  int gmpm_copy_contact_pairs_##SUF(gmpm_state_##SUF* s, size_t n, const uint32_t* mpm_id,          \
                                    const uint32_t* rigid_id, const T* dist, const T* normal,       \
                                    const T* pos, const T* rigid_v, const T* rigid_p_WB);           \
  /* Replaces DeformableDriver::CalcMpmContactPairs + CopyContactPairs (deformable_driver.h:120-194,   */\
  /* cuda_mpm_solver.cu:194-212) when every rigid geometry is an analytic box or sphere: the pairs     */\
  /* are generated on the device from the last SyncParticleStateToCpu snapshot, in the order the host  */\
  /* loop would produce them (geometry by geometry, particles in visible order), with                  */\
  /* penetration_distance = signed distance < 0 and normal = -gradient.  *n_pairs <- number of pairs.  */\
  int gmpm_generate_contact_pairs_##SUF(gmpm_state_##SUF* s, const gmpm_analytic_geometry* geoms,   \
                                        size_t n_geoms, size_t* n_pairs);                           \
  /* The whole DeformableDriver substep loop (deformable_driver.h:240-260) n times in one call, for    */\
  /* analytic rigid geometry: RebuildMapping (sorted every sort_every-th substep), FEM, P2G,          */\
  /* UpdateGrid, device-side contact query, UpdateContact, G2P.  contact_query_frequency is           */\
  /* MpmConfigParams::contact_query_frequency (cpu_mpm_model.h:23, unused upstream): pairs are        */\
  /* regenerated every that many substeps (and after every re-sort) and reused in between.  The        */\
  /* geometries move with their velocity from the given pose.  With n_geoms = 0 this is run_substeps.  */\
  int gmpm_run_driver_substeps_##SUF(gmpm_state_##SUF* s, int n, T dt, int mpm_bc, int sort_every,  \
                                     const gmpm_analytic_geometry* geoms, size_t n_geoms,           \
                                     int contact_query_frequency, T friction_mu, T stiffness,       \
                                     T damping, int exact_line_search, int max_iterations,          \
                                     gmpm_driver_stats* stats);                                     \
  /* frozen != 0: gmpm_run_driver_substeps keeps the geometries at the given pose for all n substeps */ \
  /* (the reference driver's behaviour: the rigid poses of deformable_driver.h:240-260 are those of   */\
  /* the frame's context, constant over its substeps).  Default 0: they advance by velocity*dt.       */\
  int gmpm_set_driver_geometry_frozen_##SUF(gmpm_state_##SUF* s, int frozen);                       \
  /* GpuMpmSolver<T>::UpdateContact, cuda_mpm_solver.cu:215-621.  max_iterations <= 0 -> 2000. */   \
  int gmpm_update_contact_##SUF(gmpm_state_##SUF* s, T dt, T friction_mu, T stiffness, T damping,   \
                                int exact_line_search, int max_iterations,                          \
                                gmpm_contact_stats* stats);                                         \
  /* GpuMpmState<T>::ReallocateExternelBodies, cuda_mpm_model.cu:320-338 (zeroes the buffers) */    \
  int gmpm_reallocate_external_bodies_##SUF(gmpm_state_##SUF* s, size_t n_bodies);                  \
  /* GpuMpmState<T>::ExternelBodyForceToHost, cuda_mpm_model.cu:341-345: tau[3n], f[3n] */          \
  int gmpm_external_body_force_to_host_##SUF(gmpm_state_##SUF* s, T* tau, T* f);                    \
  /* GpuMpmState<T>::DumpCpuState, cuda_mpm_model.cu:245-265: vert_pos[3*n_verts], tri[3*n_faces] */\
  int gmpm_dump_cpu_state_##SUF(gmpm_state_##SUF* s, T* vert_pos, int* tri);                        \
  /* sizes: n_verts()/n_faces()/n_particles()/num_contacts(), grid_touched_cnt_host() */            \
  size_t gmpm_n_verts_##SUF(const gmpm_state_##SUF* s);                                             \
  size_t gmpm_n_faces_##SUF(const gmpm_state_##SUF* s);                                             \
  size_t gmpm_n_particles_##SUF(const gmpm_state_##SUF* s);                                         \
  size_t gmpm_n_volumetric_##SUF(const gmpm_state_##SUF* s); /* particles added by gmpm_add_particles */\
  size_t gmpm_n_cells_##SUF(const gmpm_state_##SUF* s);                                             \
  size_t gmpm_num_contacts_##SUF(const gmpm_state_##SUF* s);                                        \
  int gmpm_grid_touched_cnt_host_##SUF(gmpm_state_##SUF* s, uint32_t* out);                         \
  /* Raw-accessor replacement (cuda_mpm_model.cuh:43-143): copies the named buffer to the host in */\
  /* the REFERENCE's layout and visible particle order.  Names = the reference accessor names     */\
  /* without the current_ prefix: positions velocities volumes affine_matrices pids sort_keys     */\
  /* sort_ids forces taus index_mappings indices Dm_inverses deformation_gradients grid_masses    */\
  /* grid_momentum grid_v_star grid_touched_flags grid_touched_ids grid_Hess grid_Grad grid_Dir   */\
  /* contact_vel contact_vel0 contact_sort_keys contact_mpm_id contact_rigid_id contact_dist        */\
  /* contact_normal contact_pos contact_rigid_v contact_rigid_p.  *bytes_written <- size; fails if  */\
  /* cap too small.                                                                                   */\
  int gmpm_read_##SUF(gmpm_state_##SUF* s, const char* name, void* out, size_t cap_bytes,           \
                      size_t* bytes_written);                                                       \
  /* ---- multi-GPU extension (not in the reference: settings.h:40 G_DEVICE_COUNT = 1 is unused). ---- */\
  /* One process and one state per GPU; every rank holds whole cloth meshes and addresses the same    */\
  /* grid.  Between ParticleToGrid and UpdateGrid the ranks sum the grid blocks more than one of them */\
  /* reaches (Multi-GPU).  The library does the device side; the host runs two all-reduces */\
  /* (NCCL SUM) per substep between the calls.  PLAN, for the next substep, off the critical path:    */\
  /*   halo_flags      : *flags_dev = uint8[*n_blocks], 1 within one block of a block this rank's P2G */\
  /*                     touched   -> host all-reduces flags_dev in place on gmpm_halo_stream         */\
  /*   halo_plan_async : shared blocks = flags > 1 in ascending block id (identical on every rank),   */\
  /*                     built on the halo stream; nothing is returned and nothing blocks             */\
  /* EXCHANGE, with the list planned one substep earlier (the first one: plan, then exchange):        */\
  /*   halo_pack       : *packed_dev = T[*n_shared][4][64] = this rank's mass and momentum of the     */\
  /*                     shared blocks (zeros where it has none)                                      */\
  /*                               -> host all-reduces packed_dev[0 : *n_shared * 256] on gmpm_stream */\
  /*   halo_unpack     : the sums are written back into the blocks this rank touched; UpdateGrid next */\
  /* halo_pack fails if a block was touched outside the set planned for it (a particle moved more     */\
  /* than one grid block in one substep).                                                             */\
  int gmpm_halo_flags_##SUF(gmpm_state_##SUF* s, void** flags_dev, size_t* n_blocks);               \
  int gmpm_halo_plan_async_##SUF(gmpm_state_##SUF* s);                                              \
  /* forgets the shared-block plan (made one substep ahead for the old particle ownership): the next  */\
  /* exchange plans afresh after its P2G, as at start-up.  Call on EVERY rank after a migration.      */\
  int gmpm_halo_reset_plan_##SUF(gmpm_state_##SUF* s);                                              \
  int gmpm_halo_pack_##SUF(gmpm_state_##SUF* s, uint32_t* n_shared, void** packed_dev);             \
  int gmpm_halo_unpack_##SUF(gmpm_state_##SUF* s);                                                  \
  /* Peer mode (<= 8 ranks of one NVSwitch domain): the exchange itself goes through receive windows  */\
  /* in peer memory -- halo_peer_send stores this rank's block values straight into the windows of    */\
  /* the ranks sharing each block and raises their delivery flags; halo_peer_recv waits for its       */\
  /* sharers ON THE DEVICE and sums in ascending rank order (identical bits on every sharer).  They    */\
  /* replace halo_pack / all-reduce / halo_unpack; the plan half (halo_flags / halo_plan_async) is     */\
  /* unchanged, its flags now carry 1 << rank.  Set-up, once, after finalize and before the first      */\
  /* halo_flags: peer_init on every rank; peer_window yields the CUDA IPC handle (64 bytes; other      */\
  /* processes) and the raw pointer (states of one process) of this rank's window; peer_attach maps    */\
  /* every other rank's window (exactly one of ipc_handle64 / same_process_ptr non-NULL).              */\
  int gmpm_halo_peer_init_##SUF(gmpm_state_##SUF* s, uint32_t rank, uint32_t world, uint32_t cap_blocks); \
  int gmpm_halo_peer_window_##SUF(gmpm_state_##SUF* s, unsigned char* ipc_handle64, void** local_ptr); \
  int gmpm_halo_peer_attach_##SUF(gmpm_state_##SUF* s, uint32_t peer, const unsigned char* ipc_handle64, \
                                  void* same_process_ptr);                                          \
  int gmpm_halo_peer_send_##SUF(gmpm_state_##SUF* s, uint32_t* n_shared);                           \
  int gmpm_halo_peer_recv_##SUF(gmpm_state_##SUF* s);                                               \
  /* UpdateGrid and GridToParticle in two parts, to hide the all-reduce of the exchange: part 0 =    */\
  /* the blocks that are not exchanged and the particles whose stencils stay clear of exchanged     */\
  /* blocks (run between halo_pack and halo_unpack, while the all-reduce is in flight), part 1 = the */\
  /* rest (after halo_unpack).  Order: update(0) g2p(0) [unpack] update(1) g2p(1); the result is     */\
  /* that of gmpm_update_grid + gmpm_g2p.                                                            */\
  int gmpm_update_grid_part_##SUF(gmpm_state_##SUF* s, int mpm_bc, int part);                       \
  int gmpm_g2p_part_##SUF(gmpm_state_##SUF* s, T dt, int part);                                     \
  /* The cudaStream_t all work of this state is issued on, and the side stream of the halo planning  */\
  /* (for stream-ordered collectives). */                                                           \
  int gmpm_stream_##SUF(gmpm_state_##SUF* s, void** cuda_stream);                                   \
  int gmpm_halo_stream_##SUF(gmpm_state_##SUF* s, void** cuda_stream);                              \
  /* Checkpoint / restore (not in the reference: its state cannot be resumed, DumpCpuState keeps vertex */\
  /* positions only, cuda_mpm_model.cu:245-265).  The blob holds positions, velocities, volumes, affine */\
  /* matrices, deformation gradients, Dm^-1 and pids in the reference's layout and visible order;       */\
  /* load_state needs a finalized state built from the same meshes.                                     */\
  size_t gmpm_checkpoint_bytes_##SUF(const gmpm_state_##SUF* s);                                    \
  int gmpm_save_state_##SUF(gmpm_state_##SUF* s, void* out, size_t cap_bytes, size_t* bytes_written);\
  int gmpm_load_state_##SUF(gmpm_state_##SUF* s, const void* in, size_t bytes);                     \
  /* One substep of the driver loop without contact (deformable_driver.h:240-260 /                */\
  /* cuda_mpm_test.cc:66-72): RebuildMapping(sort) CalcFemStateAndForce ParticleToGrid            */\
  /* UpdateGrid(mpm_bc) GridToParticle, n_substeps times; sort_every = k sorts at substeps        */\
  /* 0,k,2k.. (0 = never).  Same results as the individual calls (one ABI crossing).              */\
  int gmpm_run_substeps_##SUF(gmpm_state_##SUF* s, int n_substeps, T dt, int mpm_bc,                \
                              int sort_every);                                                      \
  /* Instrumentation: number of kernels launched so far by this state, and per-kernel device time.*/\
  uint64_t gmpm_kernel_launches_##SUF(const gmpm_state_##SUF* s);                                   \
  /* profiling on: every kernel is bracketed by CUDA events on the state's stream (graphs off).   */\
  int gmpm_set_profiling_##SUF(gmpm_state_##SUF* s, int on);                                        \
  /* total ms and launches of kernels whose name starts with `prefix` since profiling was enabled */\
  int gmpm_kernel_time_##SUF(gmpm_state_##SUF* s, const char* prefix, double* total_ms,             \
                             uint64_t* launches);                                                   \
  /* writes "name ms launches\n" lines for every profiled kernel; returns bytes needed */          \
  size_t gmpm_profile_report_##SUF(gmpm_state_##SUF* s, char* buf, size_t cap);                     \
  /* CUDA-event stopwatch on the state's stream: mark slot (0..15), then elapsed ms between two    */\
  /* marks (synchronises on the later one).  Used by bench.py to time on the device.              */\
  int gmpm_timer_mark_##SUF(gmpm_state_##SUF* s, int slot);                                         \
  int gmpm_timer_elapsed_ms_##SUF(gmpm_state_##SUF* s, int slot_begin, int slot_end, double* ms);   \
  /* device bytes currently held by the state */                                                   \
  size_t gmpm_device_bytes_##SUF(const gmpm_state_##SUF* s);                                        \
  /* binning summary of the last RebuildMapping: particle blocks and touched grid blocks */        \
  int gmpm_block_counts_##SUF(gmpm_state_##SUF* s, uint32_t* particle_blocks,                       \
                              uint32_t* touched_blocks);

GMPM_DECLARE(f32, float)
GMPM_DECLARE(f64, double)

/* Integer helpers of the path (cuda_mpm_kernels.cuh:296-363), host-callable for hosts that need
 * the same cell addressing (contact queries, I/O). */
uint32_t gmpm_morton_code(uint32_t x, uint32_t y, uint32_t z);
uint32_t gmpm_cell_index(uint32_t xi, uint32_t yi, uint32_t zi);
void gmpm_inverse_cell_index(uint32_t index, uint32_t* xyz);
/* Stable device radix sort of (key,id) pairs on key bits [0,num_bit) -- the reference's radix_sort
 * (radix_sort.cuh:202-286) semantics.  Host arrays in/out (testing / standalone use). */
int gmpm_radix_sort_pairs(const uint32_t* keys_in, const uint32_t* ids_in, size_t n, int num_bit,
                          uint32_t* keys_out, uint32_t* ids_out);

/* No counterpart in the reference (its radix_sort cudaMallocs its block sums on every call,
 * radix_sort.cuh:202-286): the same stable sort as gmpm_radix_sort_pairs, for callers that are not
 * allowed to allocate device memory.  Who that is: a host whose GPU memory was handed out once at
 * start-up and is accounted for byte by byte (a simulator embedded in a real-time control loop, where
 * cudaMalloc may take a device-wide lock and stall every other stream for milliseconds); code that
 * and cudaFree waits for the whole device to drain); and a plugin that runs inside a larger
 * application which owns the device allocator and forbids cudaMalloc/cudaFree behind its back.
 * This entry point calls no cudaMalloc, no cudaFree and no library routine that allocates.
 *
 * What differs from gmpm_radix_sort_pairs:
 *   - the four arrays are DEVICE pointers of n uint32 each on the current device, supplied and owned
 *     by the caller (with nothing allocated there is no place to stage host arrays); inputs are only
 *     read, outputs must not overlap the inputs or each other;
 *   - the work is enqueued on cuda_stream (a cudaStream_t; NULL = the default stream) and the call
 *     returns without waiting: synchronise that stream before reading the outputs;
 *   - the second ping-pong buffer and the digit histograms live in an array built into the library
 *     image, 12.6 MB (2 x 4 MB + 4.2 MB) that every GPU which loads libgmpm_b200.so carries until the
 *     process exits.  Its size fixes the LIMIT: n <= GMPM_RADIX_SORT_STATIC_MAX_PAIRS; a larger n
 *     fails with GMPM_ERR_INVALID and a message naming the limit, and nothing is launched;
 *   - there is one such array per GPU and every call on that GPU uses it.  Two calls that can run
 *     at the same time on one GPU (different streams, different host threads, different states)
 *     corrupt each other's result: keep them on one stream or order them with events.  Nothing in the
 *     array needs clearing between calls -- each pass writes what it later reads -- so back-to-back
 *     calls on one stream are fine, whatever the previous call left behind;
 *   - the first call on each device looks up the array's address on that device and the launch
 *     geometry (a few runtime queries, no allocation).  Make that first call outside any stream
 *     capture; later calls on the device skip the set-up.
 * Results are bit-identical to gmpm_radix_sort_pairs (same kernel, same pass plan). */
#define GMPM_RADIX_SORT_STATIC_MAX_PAIRS (1u << 20)
int gmpm_radix_sort_pairs_static_scratch(const uint32_t* keys_in_dev, const uint32_t* ids_in_dev, size_t n,
                                         int num_bit, uint32_t* keys_out_dev, uint32_t* ids_out_dev,
                                         void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* GMPM_C_H_ */
