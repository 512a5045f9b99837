#include "gptq_solver_super.cuh"
namespace gq {
int launch_solve_super_m0_nq3(const SuperMaps& maps_u, const SuperParams& p, cudaStream_t st) {
  return launch_super<0, 3, uint8_t>(maps_u, p, st);
}
}  // namespace gq
