// Host driver for K3 + K4 (see gptq_solver_kernel.cuh for the kernel and its design notes).
// The six kernel instantiations live in gptq_solver_inst_*.cu so they compile in parallel
// (each is ~10k fully unrolled instructions).
#include <algorithm>
#include <vector>
#include "gptq_b200_internal.h"

namespace gq {

struct SolveParams;
#define GQ_DECL(MODE, NB) int launch_solve_block_m##MODE##_nb##NB(const SolveParams& p, cudaStream_t st);
GQ_DECL(0, 32) GQ_DECL(0, 64) GQ_DECL(0, 128) GQ_DECL(1, 32) GQ_DECL(1, 64) GQ_DECL(1, 128)
#undef GQ_DECL
struct SuperParams;
struct SuperMaps;
#define GQ_DECL(MODE, NQ) int launch_solve_super_m##MODE##_nq##NQ(const SuperMaps& maps_u, const SuperParams& p, cudaStream_t st);
GQ_DECL(0, 1) GQ_DECL(0, 2) GQ_DECL(0, 3) GQ_DECL(0, 4) GQ_DECL(1, 1) GQ_DECL(1, 2) GQ_DECL(1, 3) GQ_DECL(1, 4)
#undef GQ_DECL

}  // namespace gq
#include "gptq_solver_params.h"
namespace gq {

// rows per CTA of the fused super-block kernel: the smallest tile that still fits the linear into one wave of CTAs
static int launch_super_nq(int mode, const SuperMaps& map_u, const SuperParams& p, int n_sms, cudaStream_t st) {
  const int nq = (p.R + 31) / 32 <= n_sms ? 1 : (p.R + 63) / 64 <= n_sms ? 2 : (p.R + 95) / 96 <= n_sms ? 3 : 4;
  if (mode == 0) {
    switch (nq) {
      case 1: return launch_solve_super_m0_nq1(map_u, p, st);
      case 2: return launch_solve_super_m0_nq2(map_u, p, st);
      case 3: return launch_solve_super_m0_nq3(map_u, p, st);
      default: return launch_solve_super_m0_nq4(map_u, p, st);
    }
  }
  switch (nq) {
    case 1: return launch_solve_super_m1_nq1(map_u, p, st);
    case 2: return launch_solve_super_m1_nq2(map_u, p, st);
    case 3: return launch_solve_super_m1_nq3(map_u, p, st);
    default: return launch_solve_super_m1_nq4(map_u, p, st);
  }
}

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
// mode 3: mode 1's arithmetic (static parameters, 16-bit-rounded dequant) with the per-group
// (C/group, R) parameter layout of mode 0: all groups of a linear in ONE call (one tensor-map
// build, no host round trip per group).
// mode 4: mode 0's fp32 arithmetic with caller-supplied per-group parameters (symmetric / MSE-refined / EXL2 grids).
// maxq_groups (host, C/group entries, may be null) overrides maxq per group in modes 0, 3 and 4.
int gptq_solve(float* Wt, const float* U, float* quant_t, void* codes_t, int codes_elem_size, float* err_t,
               float* scale_gr, float* qzero_gr, int R, int C, int group, float maxq, int mode, int a, int b,
               cudaStream_t st, const float* maxq_groups) {
  if (R % 4 != 0 || C % 4 != 0) return GQ_ERR_SHAPE;           // float4 path, as the reference (gptq_cpp_kernel.cu:130)
  if (a < 0 || b > C || a >= b) return GQ_ERR_ARG;
  const bool per_group = mode == 0 || mode == 3 || mode == 4;
  const int static_params = mode == 4;
  if (mode == 3) mode = 1;
  if (mode == 4) mode = 0;
  // codes are uint8 in mode 0 (this build's storage type) and int16 in mode 1 (the reference's out_q)
  if (codes_t && codes_elem_size != (mode == 0 ? 1 : 2)) return GQ_ERR_ARG;
  if (per_group && (group != 32 && group != 64 && group != 128)) return GQ_ERR_SHAPE;
  if (per_group && ((a % group) || (b % group))) return GQ_ERR_SHAPE;
  if ((a % 4) != 0) return GQ_ERR_SHAPE;                        // 16-byte aligned U row starts
  // ---- plan the blocks BEFORE the first launch: an unsupported range must not leave half-updated tensors behind.
  // Per-group modes walk `group`-sized blocks; the quantize_range contract (mode 1) takes any [a, b) whose length is a
  // multiple of 32 and is cut greedily into 128 / 64 / 32-column blocks (same updates as the reference's per-column
  // loop, applied in block order).
  std::vector<int> blocks;
  for (int c = a; c < b;) {
    int nb = group;
    if (!per_group) {
      const int left = b - c;
      nb = left >= 128 ? 128 : left >= 64 ? 64 : left >= 32 ? 32 : 0;
      if (nb == 0) return GQ_ERR_SHAPE;
    }
    blocks.push_back(nb);
    c += nb;
  }
  Tf32Maps maps;
  const bool tc = g_use_tf32x3 != 0;
  if (tc) {
    int rcm = gemm_tf32x3_maps(&maps, U, C, C, C, err_t, C, R, R, Wt, C, R, R);
    if (rcm != GQ_OK) return rcm;
  }
  // ---- fused path (gptq_solver_super.cuh): 128-column blocks only.  One launch solves a super-block of up to four blocks
  // and keeps their mutual updates in tensor memory; the columns behind it get one rank-(super-block) update.
  bool all128 = tc && g_solve_fused && g_solve_super > 1;
  for (int nb : blocks) all128 = all128 && nb == 128;
  if (all128) {
    static int n_sms = 0;
    if (n_sms == 0) {
      int dev = 0;
      cudaGetDevice(&dev);
      if (cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sms <= 0) n_sms = 148;
    }
    SuperMaps map_u;
    int rc = GQ_OK;
    for (int g = 0; g < 3; ++g)
      if ((rc = make_tmap_f32_colgroups(&map_u.m[g], U, (uint64_t)C, (uint64_t)C, (uint64_t)C, 8, 4 * (g + 1))) != GQ_OK) return rc;
    for (int g = 0; g < 8; ++g)
      if ((rc = make_tmap_2d_raw(&map_u.upk[g], U, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (uint64_t)C, (uint64_t)C, (uint64_t)C * 4,
                                 128 - 16 * g, 16, CU_TENSOR_MAP_SWIZZLE_NONE)) != GQ_OK) return rc;
    const int per = std::min(g_solve_super, (int)SUP_MAXBLK);
    const int n_blocks = (int)blocks.size();
    for (int b0 = 0; b0 < n_blocks; b0 += per) {
      SuperParams sp;
      sp.Wt = Wt; sp.U = U; sp.quant_t = quant_t; sp.codes_t = codes_t; sp.err_t = err_t;
      sp.dbg = (b0 == 0) ? g_solve_dbg : nullptr;
      sp.R = R; sp.C = C; sp.c0 = a + 128 * b0; sp.nblk = std::min(per, n_blocks - b0); sp.static_params = static_params;
      for (int i = 0; i < SUP_MAXBLK; ++i)
        sp.maxq[i] = (per_group && maxq_groups && i < sp.nblk) ? maxq_groups[sp.c0 / 128 + i] : maxq;
      if (per_group) {
        sp.scale = scale_gr + (size_t)(sp.c0 / 128) * R; sp.qzero = qzero_gr + (size_t)(sp.c0 / 128) * R; sp.param_stride = R;
      } else {
        sp.scale = scale_gr; sp.qzero = qzero_gr; sp.param_stride = 0;
      }
      if ((rc = launch_super_nq(mode, map_u, sp, n_sms, st)) != GQ_OK) return rc;
      const int e = sp.c0 + 128 * sp.nblk;
      if (e < C && (rc = gemm_tf32x3(maps, C - e, R, e - sp.c0, sp.c0, e, sp.c0, 0, e, 0, -1.f, 0, st)) != GQ_OK) return rc;
    }
    return GQ_OK;
  }
  // Two-level lazy batching: inside a super-block of `super` x 128 columns the update only runs over the super-block's
  // own later columns (they are needed next); the columns behind it get ONE rank-(super-block width) update when the
  // super-block is finished: a quarter of the read-modify-write passes over W and a deeper, more efficient GEMM.
  // The end of a super-block is fixed when it opens (blocks of different widths may share one).
  const int super_cols = (tc ? g_solve_super : 1) * 128;
  int c = a;
  int sb_start = a, sb_end = std::min(std::min(a + super_cols, b), C);
  for (size_t bi = 0; bi < blocks.size(); ++bi) {
    const int nb = blocks[bi];
    SolveParams p;
    p.Wt = Wt; p.U = U; p.quant_t = quant_t; p.codes_t = codes_t; p.err_t = err_t;
    p.R = R; p.C = C; p.c0 = c; p.maxq = maxq; p.static_params = static_params;
    if (per_group) {
      if (maxq_groups) p.maxq = maxq_groups[c / group];
      p.scale = scale_gr + (size_t)(c / group) * R;
      p.qzero = qzero_gr + (size_t)(c / group) * R;
    } else {
      p.scale = scale_gr; p.qzero = qzero_gr;
    }
    int rc = launch_block_nb(mode, p, nb, st);
    if (rc != GQ_OK) return rc;
    const int e = c + nb;
    const bool close_super = !tc || g_solve_super <= 1 || e >= sb_end;   // super = 1: the reference's schedule, one full trailing update per block
    if (e < C) {
      if (!tc) {
        GemmProblem g{U + (size_t)c * C + e, err_t + (size_t)c * R, Wt + (size_t)e * R, C - e, R, nb, C, R, R};
        if ((rc = sgemm(g, -1.f, 1.f, GEMM_TRANS_A, st)) != GQ_OK) return rc;
      } else if (close_super) {
        // everything behind the super-block at once: rows [e, C) all take U[sb_start:e, .]^T E[sb_start:e, :] (the
        // blocks before this one already reached rows [., e) through their near updates)
        if ((rc = gemm_tf32x3(maps, C - e, R, e - sb_start, sb_start, e, sb_start, 0, e, 0, -1.f, 0, st)) != GQ_OK) return rc;
      } else {
        // W[e:sb_end, :] -= U[c:e, e:sb_end]^T E[c:e, :]
        if ((rc = gemm_tf32x3(maps, sb_end - e, R, nb, c, e, c, 0, e, 0, -1.f, 0, st)) != GQ_OK) return rc;
      }
    }
    if (close_super) { sb_start = e; sb_end = std::min(std::min(e + super_cols, b), C); }
    c = e;
  }
  return GQ_OK;
}

}  // namespace gq
