// one instantiation of the block solver (mode 0, 128 columns) -- see gptq_solver_kernel.cuh
#include "gptq_solver_kernel.cuh"
namespace gq {
int launch_solve_block_m0_nb128(const SolveParams& p, cudaStream_t st) { return launch_block<0, 128, uint8_t>(p, st); }
}  // namespace gq
