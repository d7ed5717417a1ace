// C++ twin of the reference's only MPM test (multibody/gpu_mpm/cuda_mpm_test.cc:18-79) and of the
// DeformableDriver substep (multibody/plant/deformable_driver.h:221-271), written against the drop-in
// headers include/multibody/gpu_mpm/*.  Unlike the reference's test (which asserts a non-null pointer),
// every step is checked against the CPU oracle (oracle/gmpm_oracle.h; test infrastructure).
//   usage: test_shim [res] [substeps]     exit code 0 = parity ok
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "multibody/gpu_mpm/cuda_mpm_solver.cuh"

extern "C" {
#include "gmpm_oracle.h"
}

using namespace drake::multibody::gmpm;
using T = double;

static double rel(const std::vector<T>& a, const T* b) {
  double num = 0, den = 0;
  for (size_t i = 0; i < a.size(); ++i) {
    num = std::fmax(num, std::fabs((double)a[i] - (double)b[i]));
    den = std::fmax(den, std::fabs((double)b[i]));
  }
  return num / std::fmax(den, 1e-300);
}

// Compile-time coverage of the extension methods for both scalars (their run-time parity is in
// tests/test_gpu_contact.py and tests/test_gpu_parity.py, through the same C ABI entry points).
template <typename S>
void instantiate_extensions(GpuMpmState<S>* st, const GpuMpmSolver<S>& sv) {
  std::vector<gmpm_analytic_geometry> geoms;
  sv.GenerateContactPairs(st, geoms);
  sv.RunDriverSubsteps(st, 1, S(1e-3), geoms, -1, 0, 1, S(0), S(0), S(0), false);
  st->LoadState(st->SaveState());
  gmpm_material mat{GMPM_MAT_DRUCKER_PRAGER, 0, 1e5, 0.3, 30.0, 0.0};
  st->AddParticles({}, {}, {}, mat);
  std::vector<S> rec;
  std::vector<int> pids;
  const S box[6] = {0, 0, 0, 1, 1, 1};
  st->MigrateOut(0, S(0), S(1), box, &rec, &pids, 0);
  st->MigrateIn(rec);
  st->SetContactQueryBounds({});
  st->ClearContactQueryBounds();
}
template void instantiate_extensions<float>(GpuMpmState<float>*, const GpuMpmSolver<float>&);
template void instantiate_extensions<double>(GpuMpmState<double>*, const GpuMpmSolver<double>&);

int main(int argc, char** argv) {
  const int res = argc > 1 ? atoi(argv[1]) : 40;
  const int substeps = argc > 2 ? atoi(argv[2]) : 30;
  const T dt = 1e-3;
  // the sheet of cuda_mpm_test.cc:27-57, lowered so that it meets a box collider
  std::vector<Vec3<T>> pos, vel;
  std::vector<int> tri;
  const T l = T(0.5) / res;
  for (int i = 0; i < res; ++i)
    for (int j = 0; j < res; ++j) {
      pos.emplace_back(T(0.25) + i * l, T(0.25) + j * l, T(0.3) + T(0.01) * std::sin(20 * i * l));
      vel.emplace_back(T(0), T(0), T(-0.5));
    }
  auto p = [res](int i, int j) { return i * res + j; };
  for (int i = 0; i + 1 < res; ++i)
    for (int j = 0; j + 1 < res; ++j) {
      for (int v : {p(i, j), p(i + 1, j), p(i, j + 1), p(i + 1, j + 1), p(i, j + 1), p(i + 1, j)}) tri.push_back(v);
    }

  GpuMpmState<T> state;
  state.AddQRCloth(pos, vel, tri);
  state.Finalize();
  GpuMpmState<T> alias = state;  // Drake copies the state by value (deformable_model.cc:428): both see one device state
  GpuMpmSolver<T> solver;

  ogmpm_params op;
  ogmpm_default_params(&op);
  ogmpm_state_f64* orc = ogmpm_create_f64(&op);
  ogmpm_add_qr_cloth_f64(orc, reinterpret_cast<const T*>(pos.data()), reinterpret_cast<const T*>(vel.data()),
                         pos.size(), tri.data(), tri.size() / 3);
  ogmpm_finalize_f64(orc);

  const size_t n_bodies = 2;
  double worst = 0;
  long contacts_seen = 0;
  for (int s = 0; s < substeps; ++s) {
    // InitalizeExternalContactForces (deformable_driver.h:196-208)
    alias.ReallocateExternelBodies(n_bodies);
    alias.external_forces_host().resize(n_bodies);
    ogmpm_set_num_bodies_f64(orc, n_bodies);
    // the substep (deformable_driver.h:240-260)
    solver.SyncParticleStateToCpu(&alias);
    solver.RebuildMapping(&alias, s % 10 == 0);
    solver.CalcFemStateAndForce(&alias, dt);
    solver.ParticleToGrid(&alias, dt);
    solver.UpdateGrid(&alias, -1);
    ogmpm_rebuild_mapping_f64(orc, s % 10 == 0);
    ogmpm_calc_fem_f64(orc, dt);
    ogmpm_p2g_f64(orc, dt);
    ogmpm_update_grid_f64(orc, -1);
    // CalcMpmContactPairs stand-in: analytic box [0.2,0.8]^2 x [0.2,0.3]
    MpmParticleContactPairs<T> pairs;
    const auto ph = state.positions_host();  // host mirrors are shared between the copies as well
    for (size_t i = 0; i < ph.size(); ++i) {
      const T d = ph[i][2] - T(0.3);
      if (d < 0 && ph[i][2] > T(0.2))
        pairs.push_back((uint32_t)i, 1u, d, Vec3<T>(0, 0, -1), ph[i], Vec3<T>(0, 0, 0), Vec3<T>(0.5, 0.5, 0.25));
    }
    contacts_seen += (long)pairs.size();
    solver.CopyContactPairs(&alias, pairs);
    solver.UpdateContact(&alias, 0, s, dt, T(0), T(1e6), T(1e-5), false, s % 2 == 1);
    ogmpm_set_contacts_f64(orc, pairs.size(), pairs.particle_in_contact_index.data(), pairs.non_mpm_id.data(),
                           pairs.penetration_distance.data(), reinterpret_cast<const T*>(pairs.normal.data()),
                           reinterpret_cast<const T*>(pairs.particle_in_contact_position.data()),
                           reinterpret_cast<const T*>(pairs.rigid_v.data()),
                           reinterpret_cast<const T*>(pairs.rigid_p_WB.data()));
    ogmpm_contact_stats os;
    ogmpm_update_contact_f64(orc, dt, T(0), T(1e6), T(1e-5), s % 2 == 1, 0, &os);
    if (pairs.size() && solver.last_contact_stats().iterations != os.iterations) {
      printf("FAIL substep %d: Newton iterations %d vs oracle %d\n", s, solver.last_contact_stats().iterations,
             os.iterations);
      return 1;
    }
    solver.GridToParticle(&alias, dt);
    ogmpm_g2p_f64(orc, dt);
    // FinalizeExternalContactForces (deformable_driver.h:210-219)
    alias.ExternelBodyForceToHost();
    std::vector<T> tau(3 * n_bodies), f(3 * n_bodies);
    ogmpm_body_impulses_f64(orc, tau.data(), f.data());
    const auto& ext = state.external_forces_host();
    for (size_t b = 0; b < n_bodies; ++b)
      for (int d = 0; d < 3; ++d)
        worst = std::fmax(worst, std::fabs(ext.F_Bq_W_f[b][d] - f[3 * b + d]) / std::fmax(1e-12, std::fabs(f[3 * 1 + 2])));
    for (const char* name : {"positions", "velocities"})
      worst = std::fmax(worst, rel(state.Read<T>(name, 3 * state.n_particles()),
                                   static_cast<const T*>(ogmpm_view_f64(orc, name))));
  }
  solver.GpuSync();
  auto [vp, vt] = state.DumpCpuState();
  solver.Dump(state, "/tmp/gmpm_test_shim.obj");
  const bool ok = worst <= 1e-7 && contacts_seen > 0 && vp.size() == (size_t)res * res && vt.size() == tri.size() &&
                  state.n_particles() == pos.size() + tri.size() / 3;
  printf("%s: N=%zu substeps=%d contacts(total)=%ld worst rel err vs oracle %.3e, Newton iterations %d\n",
         ok ? "OK" : "FAIL", state.n_particles(), substeps, contacts_seen, worst, alias.total_contact_iteration_count);
  ogmpm_destroy_f64(orc);
  alias.Destroy();
  return ok ? 0 : 1;
}
