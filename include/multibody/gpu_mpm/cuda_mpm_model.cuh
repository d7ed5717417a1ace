// Drop-in replacement header for the reference's multibody/gpu_mpm/cuda_mpm_model.cuh:20-260.
// GpuMpmState<T> here is a value-semantic HANDLE to a state that lives inside libgmpm_b200.so
// (gmpm_c.h): copies share the same device state, exactly like the reference's shallow copies of raw
// device pointers (Drake copies the state by value into an AbstractValue, deformable_model.cc:428),
// but the device memory is released when the last copy goes away or at an explicit Destroy().
// Header-only, plain C++17 (no CUDA headers): link with -lgmpm_b200.
#pragma once

#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "gmpm_c.h"
#include "multibody/gpu_mpm/settings.h"

namespace drake {
namespace multibody {
namespace gmpm {

// cuda_mpm_model.cuh:20-35
template <typename T>
struct ExternalSpatialForce {
  std::vector<Vec3<T>> p_BoBq_B, F_Bq_W_tau, F_Bq_W_f;
  size_t size() const { return p_BoBq_B.size(); }
  void resize(size_t n) {
    p_BoBq_B.resize(n);
    F_Bq_W_tau.resize(n);
    F_Bq_W_f.resize(n);
  }
};

// positions_host() of the reference is a std::vector<Vec3<T>> filled by a blocking cudaMemcpy
// (cuda_mpm_model.cuh:116-117, cuda_mpm_solver.cu:187-191).  Here it is a read-only view of a pinned
// mirror owned by the library: SyncParticleStateToCpu only *starts* the snapshot (device gather in
// stream order + D2H on a copy stream) and the first access to positions_host() waits for it, so the
// PCIe transfer runs under the solver calls the driver issues in between
// (deformable_driver.h:243 -> :120-194).  Same members as the driver uses: operator[], size(), data().
template <typename T>
class HostPositions {
 public:
  HostPositions() = default;
  HostPositions(const Vec3<T>* p, size_t n) : p_(p), n_(n) {}
  const Vec3<T>& operator[](size_t i) const { return p_[i]; }
  const Vec3<T>* data() const { return p_; }
  const Vec3<T>* begin() const { return p_; }
  const Vec3<T>* end() const { return p_ + n_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }

 private:
  const Vec3<T>* p_ = nullptr;
  size_t n_ = 0;
};

namespace internal {
inline void Check(int status, const char* what) {
  if (status != GMPM_OK) throw std::runtime_error(std::string(what) + ": " + gmpm_last_error());
}
// Scalar -> entry points of the C ABI.
template <typename T> struct Abi;
#define GMPM_B200_ABI(SUF, T)                                                                        \
  template <> struct Abi<T> {                                                                        \
    using State = gmpm_state_##SUF;                                                                  \
    static constexpr auto create = gmpm_create_##SUF;                                                \
    static constexpr auto destroy = gmpm_destroy_##SUF;                                              \
    static constexpr auto add_qr_cloth = gmpm_add_qr_cloth_##SUF;                                    \
    static constexpr auto add_particles = gmpm_add_particles_##SUF;                                  \
    static constexpr auto migrate_out = gmpm_migrate_out_##SUF;                                      \
    static constexpr auto migrate_in = gmpm_migrate_in_##SUF;                                        \
    static constexpr auto migrate_count = gmpm_migrate_count_##SUF;                                  \
    static constexpr auto finalize = gmpm_finalize_##SUF;                                            \
    static constexpr auto rebuild_mapping = gmpm_rebuild_mapping_##SUF;                              \
    static constexpr auto calc_fem = gmpm_calc_fem_##SUF;                                            \
    static constexpr auto p2g = gmpm_p2g_##SUF;                                                      \
    static constexpr auto update_grid = gmpm_update_grid_##SUF;                                      \
    static constexpr auto g2p = gmpm_g2p_##SUF;                                                      \
    static constexpr auto sync = gmpm_sync_##SUF;                                                    \
    static constexpr auto sync_particle_state_to_cpu = gmpm_sync_particle_state_to_cpu_##SUF;        \
    static constexpr auto sync_particle_state_to_cpu_async = gmpm_sync_particle_state_to_cpu_async_##SUF; \
    static constexpr auto positions_host = gmpm_positions_host_##SUF;                                \
    static constexpr auto copy_contact_pairs = gmpm_copy_contact_pairs_##SUF;                        \
    static constexpr auto update_contact = gmpm_update_contact_##SUF;                                \
    static constexpr auto generate_contact_pairs = gmpm_generate_contact_pairs_##SUF;                \
    static constexpr auto run_driver_substeps = gmpm_run_driver_substeps_##SUF;                      \
    static constexpr auto set_driver_geometry_frozen = gmpm_set_driver_geometry_frozen_##SUF;        \
    static constexpr auto set_query_bounds = gmpm_set_query_bounds_##SUF;                            \
    static constexpr auto clear_query_bounds = gmpm_clear_query_bounds_##SUF;                        \
    static constexpr auto reallocate_external_bodies = gmpm_reallocate_external_bodies_##SUF;        \
    static constexpr auto external_body_force_to_host = gmpm_external_body_force_to_host_##SUF;      \
    static constexpr auto dump_cpu_state = gmpm_dump_cpu_state_##SUF;                                \
    static constexpr auto n_verts = gmpm_n_verts_##SUF;                                              \
    static constexpr auto n_faces = gmpm_n_faces_##SUF;                                              \
    static constexpr auto n_particles = gmpm_n_particles_##SUF;                                      \
    static constexpr auto num_contacts = gmpm_num_contacts_##SUF;                                    \
    static constexpr auto grid_touched_cnt_host = gmpm_grid_touched_cnt_host_##SUF;                  \
    static constexpr auto read = gmpm_read_##SUF;                                                    \
    static constexpr auto run_substeps = gmpm_run_substeps_##SUF;                                    \
    static constexpr auto checkpoint_bytes = gmpm_checkpoint_bytes_##SUF;                            \
    static constexpr auto save_state = gmpm_save_state_##SUF;                                        \
    static constexpr auto load_state = gmpm_load_state_##SUF;                                        \
  };
GMPM_B200_ABI(f32, float)
GMPM_B200_ABI(f64, double)
#undef GMPM_B200_ABI
}  // namespace internal

template <typename T>
struct GpuMpmState {
  using Abi = internal::Abi<T>;

  GpuMpmState() = default;
  // Non-default runtime parameters (grid resolution, material, device); must precede AddQRCloth.
  explicit GpuMpmState(const gmpm_params& params) { Create(&params); }

  size_t n_verts() const { return h_ ? Abi::n_verts(h_.get()) : 0; }
  size_t n_faces() const { return h_ ? Abi::n_faces(h_.get()) : 0; }
  size_t n_particles() const { return h_ ? Abi::n_particles(h_.get()) : 0; }
  size_t num_contacts() const { return h_ ? Abi::num_contacts(h_.get()) : 0; }
  size_t num_external_bodies() const { return num_external_bodies_; }
  uint32_t grid_touched_cnt_host() const {
    uint32_t c = 0;
    internal::Check(Abi::grid_touched_cnt_host(raw(), &c), "grid_touched_cnt_host");
    return c;
  }

  // host mirrors (cuda_mpm_model.cuh:117-118, 141-143)
  HostPositions<T> positions_host() const {
    if (!h_) return HostPositions<T>();
    const T* p = nullptr;
    internal::Check(Abi::positions_host(h_.get(), &p), "positions_host");  // waits for the pending snapshot
    return HostPositions<T>(reinterpret_cast<const Vec3<T>*>(p), p ? n_particles() : 0);
  }
  ExternalSpatialForce<T>& external_forces_host() { return host_->external_forces; }
  const ExternalSpatialForce<T>& external_forces_host() const { return host_->external_forces; }

  // cuda_mpm_model.cu:16-33
  void AddQRCloth(const std::vector<Vec3<T>>& pos, const std::vector<Vec3<T>>& vel, const std::vector<int>& indices) {
    if (!h_) Create(nullptr);
    if (pos.size() != vel.size() || indices.size() % 3) throw std::invalid_argument("AddQRCloth: inconsistent sizes");
    internal::Check(Abi::add_qr_cloth(raw(), reinterpret_cast<const T*>(pos.data()),
                                      reinterpret_cast<const T*>(vel.data()), pos.size(), indices.data(),
                                      indices.size() / 3),
                    "AddQRCloth");
  }
  // Extension: a particle-only body with a volumetric material (gmpm_material: fixed corotated,
  // StVK-Hencky, Drucker-Prager).  Generalises AddQRCloth; `volume` is the rest volume of each particle.
  void AddParticles(const std::vector<Vec3<T>>& pos, const std::vector<Vec3<T>>& vel, const std::vector<T>& volume,
                    const gmpm_material& material) {
    if (!h_) Create(nullptr);
    if (pos.size() != vel.size() || pos.size() != volume.size()) throw std::invalid_argument("AddParticles: inconsistent sizes");
    internal::Check(Abi::add_particles(raw(), reinterpret_cast<const T*>(pos.data()), reinterpret_cast<const T*>(vel.data()),
                                       volume.data(), pos.size(), &material),
                    "AddParticles");
  }
  // Extension: particle migration between the states of a position-owned decomposition; 26 scalars
  // per record (gmpm_c.h).  Returns the number of particles that left / the local ids the arrivals took.
  size_t MigrateOut(int axis, T lo, T hi, const T (&park_box)[6], std::vector<T>* records, std::vector<int>* pids, size_t cap) {
    records->resize(cap * 26);
    pids->resize(cap);
    size_t n = 0;
    internal::Check(Abi::migrate_out(raw(), axis, lo, hi, park_box, records->data(), pids->data(), cap, &n), "MigrateOut");
    records->resize(n * 26);
    pids->resize(n);
    return n;
  }
  // Dry run of MigrateOut (changes nothing): leavers per destination slab (edges.size() + 1 entries) and, in
  // *n_free, the inactive slots of this state.  Sum the leavers over the ranks before calling MigrateOut anywhere.
  std::vector<size_t> MigrateCount(int axis, T lo, T hi, const std::vector<double>& edges, size_t* n_free) {
    std::vector<size_t> leavers(edges.size() + 1);
    internal::Check(Abi::migrate_count(raw(), axis, lo, hi, edges.data(), edges.size(), leavers.data(), n_free), "MigrateCount");
    return leavers;
  }
  std::vector<int> MigrateIn(const std::vector<T>& records) {
    std::vector<int> pids(records.size() / 26);
    internal::Check(Abi::migrate_in(raw(), records.data(), pids.size(), pids.data()), "MigrateIn");
    return pids;
  }
  // cuda_mpm_model.cu:36-122: every device allocation happens here.
  void Finalize() {
    if (!h_) Create(nullptr);
    internal::Check(Abi::finalize(raw()), "Finalize");
  }
  // cuda_mpm_model.cu:125-242.  Drops this handle's reference; the device memory is freed with the last copy.
  void Destroy() {
    if (h_) {
      h_.reset();
      host_ = std::make_shared<Host>();
    }
  }

  // cuda_mpm_model.cu:245-265: (vertex positions, triangle indices) for visualisation.
  using DumpT = std::tuple<std::vector<Vec3<T>>, std::vector<int>>;
  DumpT DumpCpuState() const {
    std::vector<Vec3<T>> pos(n_verts());
    std::vector<int> tri(3 * n_faces());
    internal::Check(Abi::dump_cpu_state(raw(), reinterpret_cast<T*>(pos.data()), tri.data()), "DumpCpuState");
    return std::make_tuple(std::move(pos), std::move(tri));
  }

  // Extension: bounds (lo xyz, hi xyz per box) of the rigid geometry the host contact query tests against; the
  // async snapshot then moves only the particles inside one of them over PCIe (see gmpm_set_query_bounds).
  void SetContactQueryBounds(const std::vector<T>& lo_hi) {
    internal::Check(Abi::set_query_bounds(raw(), lo_hi.data(), lo_hi.size() / 6), "SetContactQueryBounds");
  }
  void ClearContactQueryBounds() { internal::Check(Abi::clear_query_bounds(raw()), "ClearContactQueryBounds"); }

  // cuda_mpm_model.cu:320-345
  void ReallocateExternelBodies(size_t n) {
    internal::Check(Abi::reallocate_external_bodies(raw(), n), "ReallocateExternelBodies");
    num_external_bodies_ = n;
  }
  void ExternelBodyForceToHost() {
    auto& ext = host_->external_forces;
    if (ext.size() != num_external_bodies_) ext.resize(num_external_bodies_);
    internal::Check(Abi::external_body_force_to_host(raw(), reinterpret_cast<T*>(ext.F_Bq_W_tau.data()),
                                                     reinterpret_cast<T*>(ext.F_Bq_W_f.data())),
                    "ExternelBodyForceToHost");
  }

  // Replacement of the raw device-pointer accessors (cuda_mpm_model.cuh:43-143): a host copy of the
  // named buffer in the reference's layout and particle order (names: gmpm_c.h, gmpm_read_*).
  template <typename U>
  std::vector<U> Read(const char* name, size_t count) const {
    std::vector<U> out(count);
    size_t written = 0;
    internal::Check(Abi::read(raw(), name, out.data(), count * sizeof(U), &written), name);
    out.resize(written / sizeof(U));
    return out;
  }

  // Checkpoint / restore (extension; DumpCpuState above keeps vertex positions only): the whole particle and
  // face state as one blob; LoadState needs a finalized state built from the same meshes.
  std::vector<unsigned char> SaveState() const {
    std::vector<unsigned char> blob(Abi::checkpoint_bytes(raw()));
    size_t written = 0;
    internal::Check(Abi::save_state(raw(), blob.data(), blob.size(), &written), "SaveState");
    blob.resize(written);
    return blob;
  }
  void LoadState(const std::vector<unsigned char>& blob) {
    internal::Check(Abi::load_state(raw(), blob.data(), blob.size()), "LoadState");
  }

  int total_contact_iteration_count = 0;  // cuda_mpm_model.cuh:175

  typename Abi::State* raw() const {
    if (!h_) throw std::runtime_error("GpuMpmState: no device state (call AddQRCloth / Finalize first)");
    return h_.get();
  }

 private:
  struct Host {
    ExternalSpatialForce<T> external_forces;
  };
  void Create(const gmpm_params* p) {
    typename Abi::State* s = nullptr;
    internal::Check(Abi::create(p, &s), "GpuMpmState");
    h_ = std::shared_ptr<typename Abi::State>(s, [](typename Abi::State* q) { Abi::destroy(q); });
  }
  std::shared_ptr<typename Abi::State> h_;
  std::shared_ptr<Host> host_ = std::make_shared<Host>();
  size_t num_external_bodies_ = 0;
};

}  // namespace gmpm
}  // namespace multibody
}  // namespace drake
