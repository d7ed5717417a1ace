// A C++ host of the multi-rank exchange: two GpuMpmState<double>
// "ranks" holding the two halves of one scene, coupled through the peer-window ABI of include/gmpm_c.h
// (gmpm_halo_peer_init / peer_window / peer_attach / peer_send / peer_recv) and the plan calls (gmpm_halo_flags /
// gmpm_halo_plan_async).  The ranks live in ONE process on one GPU -- the windows are attached by raw pointer, the
// flag all-reduce is a host-side sum -- so that the recipe is compiled and run without MPI or NCCL; with one process
// per GPU the same calls take the 64-byte CUDA IPC handles and an ncclAllReduce of the flags instead.
// The split run must reproduce the single-state run.     usage: test_two_rank_host     exit code 0 = ok
#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <tuple>
#include <vector>

#include "multibody/gpu_mpm/cuda_mpm_solver.cuh"

using namespace drake::multibody::gmpm;
using T = double;

#define CK(call)                                                                         \
  do {                                                                                   \
    const int st_ = (call);                                                              \
    if (st_ != 0) { std::fprintf(stderr, "%s -> %d: %s\n", #call, st_, gmpm_last_error()); std::exit(2); } \
  } while (0)

static void add_sheet(GpuMpmState<T>* st, int res, T z, T x0) {
  std::vector<Vec3<T>> pos, vel;
  std::vector<int> tri;
  const T l = T(0.2) / res;
  for (int i = 0; i < res; ++i)
    for (int j = 0; j < res; ++j) {
      pos.emplace_back(x0 + i * l, T(0.4) + j * l, z + T(0.004) * std::sin(25 * i * l));
      vel.emplace_back(T(0.1), T(0), T(-0.6));
    }
  auto p = [res](int i, int j) { return i * res + j; };
  for (int i = 0; i + 1 < res; ++i)
    for (int j = 0; j + 1 < res; ++j)
      for (int v : {p(i, j), p(i + 1, j), p(i, j + 1), p(i + 1, j + 1), p(i, j + 1), p(i + 1, j)}) tri.push_back(v);
  st->AddQRCloth(pos, vel, tri);
}

int main() {
  const int res = 24, n_sheets = 4, steps = 15;
  const T dt = 1e-3;
  GpuMpmSolver<T> sv;
  GpuMpmState<T> full, part[2];
  for (int s = 0; s < n_sheets; ++s) {      // sheets 0,1 -> rank 0; 2,3 -> rank 1 (stacked: they share grid blocks)
    add_sheet(&full, res, T(0.40) + T(0.012) * s, T(0.4));
    add_sheet(&part[s / 2], res, T(0.40) + T(0.012) * s, T(0.4));
  }
  full.Finalize();
  for (auto& p : part) p.Finalize();

  // windows: init, publish, attach (same process: by pointer)
  void* win[2];
  unsigned char handle[2][64];
  for (uint32_t r = 0; r < 2; ++r) CK(gmpm_halo_peer_init_f64(part[r].raw(), r, 2, 4096));
  for (uint32_t r = 0; r < 2; ++r) CK(gmpm_halo_peer_window_f64(part[r].raw(), handle[r], &win[r]));
  for (uint32_t r = 0; r < 2; ++r) CK(gmpm_halo_peer_attach_f64(part[r].raw(), 1 - r, nullptr, win[1 - r]));

  auto sync_all = [&] {
    for (auto& p : part) sv.GpuSync(&p);
    cudaDeviceSynchronize();
  };
  // plan: flags of both ranks, "all-reduced" (summed) on the host, then the shared-block list on each rank
  auto plan = [&] {
    void* flags[2];
    size_t n_blocks = 0;
    for (int r = 0; r < 2; ++r) CK(gmpm_halo_flags_f64(part[r].raw(), &flags[r], &n_blocks));
    sync_all();
    std::vector<unsigned char> a(n_blocks), b(n_blocks);
    cudaMemcpy(a.data(), flags[0], n_blocks, cudaMemcpyDeviceToHost);
    cudaMemcpy(b.data(), flags[1], n_blocks, cudaMemcpyDeviceToHost);
    for (size_t i = 0; i < n_blocks; ++i) a[i] = (unsigned char)(a[i] + b[i]);
    cudaMemcpy(flags[0], a.data(), n_blocks, cudaMemcpyHostToDevice);
    cudaMemcpy(flags[1], a.data(), n_blocks, cudaMemcpyHostToDevice);
    for (int r = 0; r < 2; ++r) CK(gmpm_halo_plan_async_f64(part[r].raw()));
  };

  uint32_t shared_seen = 0;
  bool planned = false;
  for (int it = 0; it < steps; ++it) {
    const bool sort = it == 6;
    sv.RebuildMapping(&full, sort);
    sv.CalcFemStateAndForce(&full, dt);
    sv.ParticleToGrid(&full, dt);
    sv.UpdateGrid(&full, -1);
    sv.GridToParticle(&full, dt);
    for (auto& p : part) {
      sv.RebuildMapping(&p, sort);
      sv.CalcFemStateAndForce(&p, dt);
      sv.ParticleToGrid(&p, dt);
    }
    if (!planned) { plan(); planned = true; }
    uint32_t n_shared = 0;
    for (auto& p : part) CK(gmpm_halo_peer_send_f64(p.raw(), &n_shared));   // every send is enqueued ...
    sync_all();                                                             // ... before any receive starts to wait
    for (auto& p : part) CK(gmpm_halo_peer_recv_f64(p.raw()));
    plan();                                                                  // for the next substep
    shared_seen += n_shared;
    for (auto& p : part) {
      sv.UpdateGrid(&p, -1);
      sv.GridToParticle(&p, dt);
    }
  }
  const auto ref = std::get<0>(full.DumpCpuState());
  double worst = 0, scale = 0;
  size_t off = 0;
  for (auto& p : part) {
    const auto got = std::get<0>(p.DumpCpuState());
    for (size_t i = 0; i < got.size(); ++i)
      for (int d = 0; d < 3; ++d) {
        worst = std::fmax(worst, std::fabs((double)got[i][d] - (double)ref[off + i][d]));
        scale = std::fmax(scale, std::fabs((double)ref[off + i][d]));
      }
    off += got.size();
  }
  const bool ok = shared_seen > 0 && worst <= 1e-10 * scale;
  std::printf("%s two-rank C++ host: %zu vertices, %u shared blocks exchanged in total, max |dx| %.3e (scale %.2f)\n",
              ok ? "OK" : "FAIL", off, shared_seen, worst, scale);
  full.Destroy();
  for (auto& p : part) p.Destroy();
  return ok ? 0 : 1;
}
