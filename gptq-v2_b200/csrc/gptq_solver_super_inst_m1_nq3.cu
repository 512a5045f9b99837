#include "gptq_solver_super.cuh"
namespace gq {
int launch_solve_super_m1_nq3(const SuperMaps& maps_u, const SuperParams& p, cudaStream_t st) {
  return launch_super<1, 3, int16_t>(maps_u, p, st);
}
}  // namespace gq
