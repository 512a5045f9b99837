// Host driver for K3 + K4 (see gptq_solver_kernel.cuh for the kernel and its design notes).
// The six kernel instantiations live in gptq_solver_inst_*.cu so they compile in parallel
// (each is ~10k fully unrolled instructions).
#include "gptq_b200_internal.h"

namespace gq {

struct SolveParams;
#define GQ_DECL(MODE, NB) int launch_solve_block_m##MODE##_nb##NB(const SolveParams& p, cudaStream_t st);
GQ_DECL(0, 32) GQ_DECL(0, 64) GQ_DECL(0, 128) GQ_DECL(1, 32) GQ_DECL(1, 64) GQ_DECL(1, 128)
#undef GQ_DECL

}  // namespace gq
#include "gptq_solver_params.h"
namespace gq {

static int launch_block_nb(int mode, const SolveParams& p, int nb, cudaStream_t st) {
  if (mode == 0) {
    switch (nb) {
      case 128: return launch_solve_block_m0_nb128(p, st);
      case 64: return launch_solve_block_m0_nb64(p, st);
      case 32: return launch_solve_block_m0_nb32(p, st);
    }
  } else {
    switch (nb) {
      case 128: return launch_solve_block_m1_nb128(p, st);
      case 64: return launch_solve_block_m1_nb64(p, st);
      case 32: return launch_solve_block_m1_nb32(p, st);
    }
  }
  return GQ_ERR_SHAPE;
}

// Solve columns [a, b) of the transposed weight matrix and push the lazy-batch update to
// every later column.  mode 0: `group`-sized blocks, scale_gr/qzero_gr are (C/group, R)
// outputs.  mode 1: scale_gr/qzero_gr are (R) inputs valid for the whole range (the
// reference's gptq_quantize_range contract); the range is processed in chunks of <=128.
int gptq_solve(float* Wt, const float* U, float* quant_t, void* codes_t, int codes_elem_size, float* err_t,
               float* scale_gr, float* qzero_gr, int R, int C, int group, float maxq, int mode, int a, int b,
               cudaStream_t st) {
  if (R % 4 != 0 || C % 4 != 0) return GQ_ERR_SHAPE;           // float4 path, as the reference (gptq_cpp_kernel.cu:130)
  if (a < 0 || b > C || a >= b) return GQ_ERR_ARG;
  // codes are uint8 in mode 0 (this build's storage type) and int16 in mode 1 (the reference's out_q)
  if (codes_t && codes_elem_size != (mode == 0 ? 1 : 2)) return GQ_ERR_ARG;
  if (mode == 0 && (group != 32 && group != 64 && group != 128)) return GQ_ERR_SHAPE;
  if (mode == 0 && ((a % group) || (b % group))) return GQ_ERR_SHAPE;
  if ((a % 4) != 0) return GQ_ERR_SHAPE;                        // 16-byte aligned U row starts
  Tf32Maps maps;
  const bool tc = g_use_tf32x3 != 0;
  if (tc) {
    int rcm = gemm_tf32x3_maps(&maps, U, C, C, C, err_t, C, R, R, Wt, C, R, R);
    if (rcm != GQ_OK) return rcm;
  }
  int c = a;
  while (c < b) {
    int nb = (mode == 0) ? group : ((b - c >= 128) ? 128 : (b - c));
    SolveParams p;
    p.Wt = Wt; p.U = U; p.quant_t = quant_t; p.codes_t = codes_t; p.err_t = err_t;
    p.R = R; p.C = C; p.c0 = c; p.maxq = maxq;
    if (mode == 0) {
      p.scale = scale_gr + (size_t)(c / group) * R;
      p.qzero = qzero_gr + (size_t)(c / group) * R;
    } else {
      p.scale = scale_gr; p.qzero = qzero_gr;
    }
    int rc = launch_block_nb(mode, p, nb, st);
    if (rc != GQ_OK) return rc;
    const int e = c + nb;
    if (e < C) {
      // W[e:, :] -= U[c:e, e:]^T * E[c:e, :]
      if (tc && nb % 16 == 0) {
        if ((rc = gemm_tf32x3(maps, C - e, R, nb, c, e, c, 0, e, 0, -1.f, 0, st)) != GQ_OK) return rc;
      } else {
        GemmProblem g{U + (size_t)c * C + e, err_t + (size_t)c * R, Wt + (size_t)e * R, C - e, R, nb, C, R, R};
        if ((rc = sgemm(g, -1.f, 1.f, GEMM_TRANS_A, st)) != GQ_OK) return rc;
      }
    }
    c = e;
  }
  return GQ_OK;
}

}  // namespace gq
