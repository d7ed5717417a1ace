// Drop-in replacement header for the reference's multibody/gpu_mpm/cuda_mpm_solver.cuh:19-35:
// the stateless launch front end, same method names / argument meaning / const-ness, forwarding to
// libgmpm_b200.so through the C ABI (gmpm_c.h).  Errors surface as std::runtime_error (the
// reference's CUDA_SAFE_CALL throws the same type when DEBUG=1, settings.h:11-25).
#pragma once

#include <fstream>
#include <string>

#include "multibody/gpu_mpm/cpu_mpm_model.h"
#include "multibody/gpu_mpm/cuda_mpm_model.cuh"
#include "multibody/gpu_mpm/settings.h"

namespace drake {
namespace multibody {
namespace gmpm {

template <typename T>
class GpuMpmSolver {
  using Abi = internal::Abi<T>;

 public:
  void RebuildMapping(GpuMpmState<T>* state, bool sort) const {
    internal::Check(Abi::rebuild_mapping(state->raw(), sort ? 1 : 0), "RebuildMapping");
  }
  void CalcFemStateAndForce(GpuMpmState<T>* state, const T& dt) const {
    internal::Check(Abi::calc_fem(state->raw(), dt), "CalcFemStateAndForce");
  }
  void ParticleToGrid(GpuMpmState<T>* state, const T& dt) const {
    internal::Check(Abi::p2g(state->raw(), dt), "ParticleToGrid");
  }
  void UpdateGrid(GpuMpmState<T>* state, int mpm_bc = -1) const {
    internal::Check(Abi::update_grid(state->raw(), mpm_bc), "UpdateGrid");
  }
  void GridToParticle(GpuMpmState<T>* state, const T& dt) const {
    internal::Check(Abi::g2p(state->raw(), dt), "GridToParticle");
  }
  // cuda_mpm_solver.cu:164-166: device-wide barrier.  Work is issued on per-state streams, so the
  // one-argument form (this state's stream only) is the cheaper barrier.
  void GpuSync() const { internal::Check(gmpm_device_sync(), "GpuSync"); }
  void GpuSync(GpuMpmState<T>* state) const { internal::Check(Abi::sync(state->raw()), "GpuSync"); }

  // cuda_mpm_solver.cu:187-191: positions_host() <- current particle positions.  The snapshot is taken
  // in stream order now; the copy to the host overlaps the following solver calls and the first
  // positions_host() access waits for it (cuda_mpm_model.cuh, HostPositions).
  void SyncParticleStateToCpu(GpuMpmState<T>* state) const {
    internal::Check(Abi::sync_particle_state_to_cpu_async(state->raw()), "SyncParticleStateToCpu");
  }

  // cuda_mpm_solver.cu:169-183: Wavefront OBJ of the cloth.
  void Dump(const GpuMpmState<T>& state, std::string filename) const {
    const auto [pos, tri] = state.DumpCpuState();
    std::ofstream obj(filename);
    for (const auto& p : pos) obj << "v " << p[0] << " " << p[1] << " " << p[2] << "\n";
    for (size_t f = 0; f + 2 < tri.size(); f += 3)
      obj << "f " << tri[f] + 1 << " " << tri[f + 1] + 1 << " " << tri[f + 2] + 1 << "\n";
  }

  // cuda_mpm_solver.cu:194-212
  void CopyContactPairs(GpuMpmState<T>* state, const MpmParticleContactPairs<T>& pairs) const {
    auto flat = [](const std::vector<Vec3<T>>& v) { return reinterpret_cast<const T*>(v.data()); };
    internal::Check(Abi::copy_contact_pairs(state->raw(), pairs.size(), pairs.particle_in_contact_index.data(),
                                            pairs.non_mpm_id.data(), pairs.penetration_distance.data(),
                                            flat(pairs.normal), flat(pairs.particle_in_contact_position),
                                            flat(pairs.rigid_v), flat(pairs.rigid_p_WB)),
                    "CopyContactPairs");
  }

  // Extension: CalcMpmContactPairs + CopyContactPairs on the device, for rigid geometry that is an
  // analytic box or sphere (deformable_driver.h:120-194 does this on the host with one signed-distance query per
  // particle and geometry).  Uses the positions of the last SyncParticleStateToCpu.  Returns the number of pairs.
  size_t GenerateContactPairs(GpuMpmState<T>* state, const std::vector<gmpm_analytic_geometry>& geometries) const {
    size_t n = 0;
    internal::Check(Abi::generate_contact_pairs(state->raw(), geometries.data(), geometries.size(), &n),
                    "GenerateContactPairs");
    return n;
  }

  // Extension: n substeps of the DeformableDriver loop (deformable_driver.h:240-260) in one call,
  // contact pairs regenerated on the device every params.contact_query_frequency substeps
  // (cpu_mpm_model.h:23: declared upstream, never read) while the geometries move with their velocity.
  gmpm_driver_stats RunDriverSubsteps(GpuMpmState<T>* state, int n, const T& dt,
                                      const std::vector<gmpm_analytic_geometry>& geometries, int mpm_bc, int sort_every,
                                      int contact_query_frequency, const T& friction_mu, const T& stiffness,
                                      const T& damping, bool exact_line_search, bool freeze_geometry = false) const {
    gmpm_driver_stats stats{};
    // freeze_geometry = true is the reference driver: rigid poses constant over the substeps of a frame
    internal::Check(Abi::set_driver_geometry_frozen(state->raw(), freeze_geometry ? 1 : 0), "RunDriverSubsteps");
    internal::Check(Abi::run_driver_substeps(state->raw(), n, dt, mpm_bc, sort_every, geometries.data(),
                                             geometries.size(), contact_query_frequency, friction_mu, stiffness, damping,
                                             exact_line_search ? 1 : 0, 0, &stats),
                    "RunDriverSubsteps");
    state->total_contact_iteration_count += stats.newton_iterations;
    return stats;
  }

  // cuda_mpm_solver.cu:215-621.  `dump` (per-iteration JSON to a developer path) is not reproduced.
  void UpdateContact(GpuMpmState<T>* state, const int frame, const int substep, const T& dt, const T& friction_mu,
                     const T& stiffness, const T& damping, const bool dump, const bool exact_line_search) const {
    (void)frame;
    (void)substep;
    (void)dump;
    gmpm_contact_stats stats{};
    internal::Check(Abi::update_contact(state->raw(), dt, friction_mu, stiffness, damping, exact_line_search ? 1 : 0,
                                        0, &stats),
                    "UpdateContact");
    state->total_contact_iteration_count += stats.iterations;
    last_contact_stats_ = stats;
  }
  const gmpm_contact_stats& last_contact_stats() const { return last_contact_stats_; }

  // Extension: n no-contact substeps in one call (cuda_mpm_test.cc:66-72 loop body).
  void RunSubsteps(GpuMpmState<T>* state, int n, const T& dt, int mpm_bc = -1, int sort_every = 0) const {
    internal::Check(Abi::run_substeps(state->raw(), n, dt, mpm_bc, sort_every), "RunSubsteps");
  }

 private:
  mutable gmpm_contact_stats last_contact_stats_{};
};

}  // namespace gmpm
}  // namespace multibody
}  // namespace drake
