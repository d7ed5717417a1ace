// Drop-in replacement header for the reference's multibody/gpu_mpm/cpu_mpm_model.h:16-114: the
// host-side PODs DeformableModel / DeformableDriver / the visualizers exchange with the MPM path.
// Same type names, member names and defaults; no device code.
#pragma once

#include <cstdint>
#include <vector>

#include "multibody/gpu_mpm/settings.h"

namespace drake {
namespace multibody {
namespace gmpm {

// cpu_mpm_model.h:17-26
template <typename T = config::GpuT>
struct MpmConfigParams {
  T substep_dt{T(1e-3)};
  bool write_files{false};
  T contact_stiffness{T(1e5)};
  T contact_damping{T(0)};
  T contact_friction_mu{T(0)};
  int contact_query_frequency{1};
  int mpm_bc{-1};
  bool exact_line_search{false};
};

// cpu_mpm_model.h:33-40: initial cloth + config, consumed by DeformableModel::DoDeclareSystemResources.
template <typename T>
struct CpuMpmModel {
  std::vector<Vec3<T>> cloth_pos, cloth_vel;
  std::vector<int> cloth_indices;
  MpmConfigParams<T> config;
};

// cpu_mpm_model.h:46-49: what the "mpm" abstract port carries to the visualizers.
template <typename T>
struct MpmPortData {
  std::vector<Vec3<T>> pos;
  std::vector<int> indices;
};

// cpu_mpm_model.h:74-114: SoA of (particle, rigid geometry) pairs with negative signed distance.
template <typename T>
struct MpmParticleContactPairs {
  std::vector<uint32_t> particle_in_contact_index, non_mpm_id;
  std::vector<T> penetration_distance;
  std::vector<Vec3<T>> normal, particle_in_contact_position, rigid_v, rigid_p_WB;

  void clear() {
    particle_in_contact_index.clear();
    non_mpm_id.clear();
    penetration_distance.clear();
    for (auto* v : {&normal, &particle_in_contact_position, &rigid_v, &rigid_p_WB}) v->clear();
  }
  void push_back(uint32_t particle, uint32_t body, T distance, const Vec3<T>& n, const Vec3<T>& position,
                 const Vec3<T>& body_velocity, const Vec3<T>& body_origin) {
    particle_in_contact_index.push_back(particle);
    non_mpm_id.push_back(body);
    penetration_distance.push_back(distance);
    normal.push_back(n);
    particle_in_contact_position.push_back(position);
    rigid_v.push_back(body_velocity);
    rigid_p_WB.push_back(body_origin);
  }
  size_t size() const { return non_mpm_id.size(); }
};

}  // namespace gmpm
}  // namespace multibody
}  // namespace drake
