// one instantiation of the block solver (mode 1, 32 columns) -- see gptq_solver_kernel.cuh
#include "gptq_solver_kernel.cuh"
namespace gq {
int launch_solve_block_m1_nb32(const SolveParams& p, cudaStream_t st) { return launch_block<1, 32, int16_t>(p, st); }
}  // namespace gq
