// Wrapper TU: compiles the reference's blelloch_scan.cu where it lies.  TEST INFRASTRUCTURE ONLY.
#include "multibody/gpu_mpm/blelloch_scan.cu"
