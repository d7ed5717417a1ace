// Development probe: cost of a grid-wide barrier on this GPU -- cooperative_groups grid.sync() against a hand-rolled
// sense-free counter barrier (one atomic arrive per CTA, acquire spin on a monotonically increasing counter).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/barrier_probe tools/barrier_probe.cu && ./barrier_probe
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void light_barrier(unsigned* counter, unsigned& phase) {
  __syncthreads();
  if (threadIdx.x == 0) {
    ++phase;
    __threadfence();
    atomicAdd(counter, 1u);
    const unsigned target = phase * gridDim.x;
    while (ld_acquire(counter) < target) {}
    __threadfence();
  }
  __syncthreads();
}

template <int MODE>
__global__ void probe(int iters, unsigned* counter, unsigned* sink) {
  cg::grid_group grid = cg::this_grid();
  unsigned phase = 0, acc = 0;
  for (int i = 0; i < iters; ++i) {
    acc += sink[(blockIdx.x * 7 + i) & 1023];   // a dependent global load per phase, as a real phase would have
    if (MODE == 0) grid.sync(); else light_barrier(counter, phase);
  }
  if (acc == 0xFFFFFFFFu) sink[0] = acc;
}

int main() {
  unsigned *counter, *sink;
  cudaMalloc(&counter, 4);
  cudaMalloc(&sink, 4096);
  cudaMemset(sink, 0, 4096);
  int iters = 200;
  for (int ctas : {8, 64, 148, 246, 296}) {
    for (int mode = 0; mode < 2; ++mode) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaMemset(counter, 0, 4);
        cudaEvent_t a, b;
        cudaEventCreate(&a); cudaEventCreate(&b);
        void* args[] = {&iters, &counter, &sink};
        cudaEventRecord(a);
        cudaLaunchCooperativeKernel(mode == 0 ? (void*)probe<0> : (void*)probe<1>, dim3(ctas), dim3(256), args, 0, 0);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      printf("%3d CTAs x 256 threads, %s: %.2f us per barrier phase (%s)\n", ctas, mode == 0 ? "cg grid.sync " : "light barrier",
             best * 1e3f / iters, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
