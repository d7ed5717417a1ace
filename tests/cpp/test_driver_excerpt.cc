// Drop-in compile proof (Missing 5): the BODIES of DeformableDriver's MPM methods --
// InitalizeExternalContactForces, FinalizeExternalContactForces and CalcAbstractStates,
// multibody/plant/deformable_driver.h:196-271 -- are taken VERBATIM from the reference tree at build time
// (tests/cpp/build_cpp.py extracts the lines into _build/driver_excerpt.inc; nothing is copied into this
// repository) and compiled, unchanged, against the drop-in headers include/multibody/gpu_mpm/* plus a minimal
// stand-in for the Drake types they touch (Context, State, plant, model).  CalcMpmContactPairs (:120-194) is
// Drake's geometry query (SignedDistanceToPoint, Jacobians): it is replaced by a stand-in that reads the same
// accessors (positions_host(), external_forces_host(), n_particles()) against one analytic box.
// At run time two frames of CalcAbstractStates are compared with the same solver calls issued directly.
//   usage: test_driver_excerpt      exit code 0 = ok
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <tuple>
#include <vector>

#include "multibody/gpu_mpm/cuda_mpm_solver.cuh"

namespace drake {
template <typename T> using Vector3 = multibody::gmpm::Vec3<T>;

namespace systems {
template <typename T>
class Context {
 public:
  T get_time() const { return time_; }
  T time_{0};
};
template <typename T>
class State {
 public:
  template <typename V> V& get_mutable_abstract_state(int) { return *static_cast<V*>(slot_); }
  void* slot_{nullptr};
};
}  // namespace systems

namespace multibody {
struct BodyIndex {
  explicit BodyIndex(size_t i) : i_(i) {}
  size_t i_;
};
namespace internal {

template <typename T>
struct FakePose {
  Vector3<T> p;
  const Vector3<T>& translation() const { return p; }
};
template <typename T>
struct FakePlant {
  int num_bodies() const { return 2; }
  T time_step() const { return T(4e-3); }
  BodyIndex get_body(BodyIndex i) const { return i; }
  FakePose<T> EvalBodyPoseInWorld(const systems::Context<T>&, BodyIndex i) const {
    return FakePose<T>{Vector3<T>(T(0.5), T(0.5), i.i_ ? T(0.45) : T(0))};
  }
};
template <typename T>
struct FakeManager {
  FakePlant<T> plant_;
  const FakePlant<T>& plant() const { return plant_; }
};
template <typename T>
struct FakeModel {
  gmpm::CpuMpmModel<gmpm::config::GpuT> cpu_;
  bool ExistsMpmModel() const { return true; }
  int gpu_mpm_state_index() const { return 0; }
  const gmpm::CpuMpmModel<gmpm::config::GpuT>& cpu_mpm_model() const { return cpu_; }
};

template <typename T>
class DeformableDriver {
 public:
  FakeManager<T> manager_storage_;
  FakeModel<T> model_storage_;
  const FakeManager<T>* manager_{&manager_storage_};
  const FakeModel<T>* deformable_model_{&model_storage_};
  gmpm::GpuMpmSolver<gmpm::config::GpuT> mpm_solver_;

  // stand-in for deformable_driver.h:120-194 (Drake geometry query): same accessors, one analytic box (body 1)
  void CalcMpmContactPairs(const systems::Context<T>&, gmpm::GpuMpmState<gmpm::config::GpuT>* mpm_state,
                           gmpm::MpmParticleContactPairs<gmpm::config::GpuT>* result) const {
    using GpuT = gmpm::config::GpuT;
    result->clear();
    const GpuT lo[3] = {0.3f, 0.3f, 0.40f}, hi[3] = {0.7f, 0.7f, 0.502f};
    for (size_t p = 0; p < mpm_state->n_particles(); ++p) {
      const auto x = mpm_state->positions_host()[p];
      bool in = true;
      GpuT best = 1e30f;
      int face = 0;
      for (int d = 0; d < 3; ++d) {
        in = in && x[d] > lo[d] && x[d] < hi[d];
        if (x[d] - lo[d] < best) { best = x[d] - lo[d]; face = d; }
        if (hi[d] - x[d] < best) { best = hi[d] - x[d]; face = 3 + d; }
      }
      if (!in) continue;
      gmpm::Vec3<GpuT> n(0, 0, 0);
      n[face % 3] = face < 3 ? GpuT(1) : GpuT(-1);   // -grad_W
      result->push_back(uint32_t(p), 1u, -best, n, x, gmpm::Vec3<GpuT>(0, 0, 0),
                        mpm_state->external_forces_host().p_BoBq_B[1]);
    }
  }

#include "driver_excerpt.inc"
};

}  // namespace internal
}  // namespace multibody
}  // namespace drake

using namespace drake::multibody;
using GpuT = gmpm::config::GpuT;

static void build_sheet(gmpm::GpuMpmState<GpuT>* st) {
  const int res = 40;
  std::vector<gmpm::Vec3<GpuT>> pos, vel;
  std::vector<int> tri;
  const GpuT l = GpuT(0.3) / res;
  for (int i = 0; i < res; ++i)
    for (int j = 0; j < res; ++j) {
      pos.emplace_back(GpuT(0.35) + i * l, GpuT(0.35) + j * l, GpuT(0.5) + GpuT(0.012) * std::sin(11 * (0.35f + i * l)));
      vel.emplace_back(GpuT(0), GpuT(0), GpuT(-0.5));
    }
  auto p = [res](int i, int j) { return i * res + j; };
  for (int i = 0; i + 1 < res; ++i)
    for (int j = 0; j + 1 < res; ++j)
      for (int v : {p(i, j), p(i + 1, j), p(i, j + 1), p(i + 1, j + 1), p(i, j + 1), p(i + 1, j)}) tri.push_back(v);
  st->AddQRCloth(pos, vel, tri);
  st->Finalize();
}

int main() {
  internal::DeformableDriver<double> driver;
  driver.model_storage_.cpu_.config.substep_dt = GpuT(1e-3);
  driver.model_storage_.cpu_.config.contact_stiffness = GpuT(1e5);
  driver.model_storage_.cpu_.config.contact_friction_mu = GpuT(0);   // frictionless: the solve has no order-dependent ties
  gmpm::GpuMpmState<GpuT> a, b;
  build_sheet(&a);
  build_sheet(&b);
  drake::systems::Context<double> context;
  drake::systems::State<double> state;
  state.slot_ = &a;
  // (1) the reference's own method bodies, two frames of four substeps each
  for (int frame = 0; frame < 2; ++frame) {
    context.time_ = frame * 4e-3;
    driver.CalcAbstractStates(context, &state);
  }
  // (2) the same calls issued directly
  gmpm::GpuMpmSolver<GpuT> sv;
  gmpm::MpmParticleContactPairs<GpuT> pairs;
  for (int frame = 0; frame < 2; ++frame) {
    driver.InitalizeExternalContactForces(context, &b);
    for (int s = 0; s < 4; ++s) {
      sv.SyncParticleStateToCpu(&b);
      sv.RebuildMapping(&b, false);
      sv.CalcFemStateAndForce(&b, GpuT(1e-3));
      sv.ParticleToGrid(&b, GpuT(1e-3));
      sv.UpdateGrid(&b, -1);
      driver.CalcMpmContactPairs(context, &b, &pairs);
      sv.CopyContactPairs(&b, pairs);
      sv.UpdateContact(&b, frame, s, GpuT(1e-3), GpuT(0), GpuT(1e5), GpuT(0), false, false);
      sv.GridToParticle(&b, GpuT(1e-3));
    }
    driver.FinalizeExternalContactForces(&b, GpuT(4e-3));
  }
  const auto pa = a.DumpCpuState();
  const auto pb = b.DumpCpuState();
  double worst = 0, scale = 0;
  for (size_t i = 0; i < std::get<0>(pa).size(); ++i)
    for (int d = 0; d < 3; ++d) {
      worst = std::fmax(worst, std::fabs((double)std::get<0>(pa)[i][d] - (double)std::get<0>(pb)[i][d]));
      scale = std::fmax(scale, std::fabs((double)std::get<0>(pb)[i][d]));
    }
  const double fa = a.external_forces_host().F_Bq_W_f[1][2], fb = b.external_forces_host().F_Bq_W_f[1][2];
  std::printf("%s driver excerpt: %zu vertices, max |dx| %.3e (scale %.2f), body force z %.6g vs %.6g, contacts %zu\n",
              (worst <= 1e-4 * scale && std::fabs(fa - fb) <= 5e-3 * std::fabs(fb) + 1e-6 && pairs.size() > 0) ? "OK" : "FAIL",
              std::get<0>(pa).size(), worst, scale, fa, fb, pairs.size());
  const bool ok = worst <= 1e-4 * scale && std::fabs(fa - fb) <= 5e-3 * std::fabs(fb) + 1e-6 && pairs.size() > 0;
  a.Destroy();
  b.Destroy();
  return ok ? 0 : 1;
}
