// contact.cuh -- MPM-rigid contact solve (replaces GpuMpmSolver<T>::UpdateContact and kernels K11-K19).
#pragma once
#include "state.cuh"

namespace gmpm {

template <typename T>
void State<T>::ensure_contact_grids() {
  if (g_Hess) return;
  g_Hess = dalloc<T>((size_t)k.n_cells * 9);
  g_Grad = dalloc<T>((size_t)k.n_cells * 3);
  g_Dir = dalloc<T>((size_t)k.n_cells * 3);
}

template <typename T>
void State<T>::copy_contact_pairs(size_t, const uint32_t*, const uint32_t*, const T*, const T*, const T*, const T*,
                                  const T*) {
  throw std::invalid_argument("contact solve not built yet");
}
template <typename T>
void State<T>::update_contact(T, T, T, T, bool, int, gmpm_contact_stats*) {
  throw std::invalid_argument("contact solve not built yet");
}
template <typename T>
void State<T>::reallocate_external_bodies(size_t) {
  throw std::invalid_argument("contact solve not built yet");
}
template <typename T>
void State<T>::external_body_force_to_host(T*, T*) {
  throw std::invalid_argument("contact solve not built yet");
}

}  // namespace gmpm
