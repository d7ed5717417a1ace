// Wrapper TU: compiles the reference's cuda_mpm_solver.cu where it lies (read-only, /root/reference)
// and adds the fp64 instantiation of the reference's own templates.  TEST INFRASTRUCTURE ONLY.
#include "multibody/gpu_mpm/cuda_mpm_solver.cu"
namespace drake { namespace multibody { namespace gmpm {
template class GpuMpmSolver<double>;
}}}
