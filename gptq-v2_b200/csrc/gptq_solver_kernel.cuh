#pragma once
// K3 + K4  blocked GPTQ column solver on the transposed (C, R) layout.
//
// Replaces, per block of <=128 columns, the reference's 2*128 per-column launches
// (fused_quantize_adjust_kernel + vv_mul_sub_kernel, package/gptq/gptq_cpp_kernel.cu:21-158,
// driven by quantize_range, gptq_cpp.cpp:33-84) by ONE launch, and restates the
// fp32 PyTorch recurrence (gptq_py.py:158-193) the parity tests are judged against.
//
// A CTA of 4 warps owns 32 weight rows (lane = row; rows are independent, so there is no
// cross-row communication).  The block's weights sit in shared memory as ws[column][lane]
// (conflict-free), the NB x NB diagonal block of U (= chol(H^-1), upper) is staged with
// cp.async.bulk row copies (upper-triangular part only) and read as warp-wide broadcasts.
// Per 16-column sub-block warp 0 runs the serial column recurrence on registers (compile-time
// indices, ~1 K instructions re-used by the rolled sub-block loop, so it stays in the
// instruction cache), then all 4 warps apply the rank-16 update to the remaining columns.
// All accesses to W/quant/codes/err are coalesced because consecutive lanes own consecutive
// rows of the (C, R) layout.
//
// MODE 0 ("pure", gptq_py.py:169-193): scale/zero computed at the block start from the
//   *updated* weights (quant.py:114-118), fp32 dequant, update = w - (e*u) with two
//   roundings (torch: matmul then subtract).  Bit-exact inside a block.
// MODE 1 ("kernel path", gptq_cpp_kernel.cu:41-73,113-137): caller-supplied static
//   scale/zero, dequant rounded through fp16, update = fma(-e, u, w).
//
// The lazy-batch trailing update  W[b:, :] -= U[a:b, b:]^T E[a:b, :]  runs after each block
// as an fp32 FMA GEMM (sgemm_f32.cu), as torch.addmm_ does in the reference with TF32 off.
#include <cuda_fp16.h>
#include "gptq_b200_internal.h"
#include "ptx.cuh"
#include "gptq_solver_params.h"

namespace gq {


constexpr int SB = 16;          // columns per register sub-block
constexpr int ROWS = 32;        // weight rows per CTA (one per lane)
constexpr int SOLVE_WARPS = 4;

// Correctly rounded a / b from rb = RN(1/b) in three dependent operations (Markstein): q = a*rb,
// r = fma(-q, b, a) (exact), q' = fma(r, rb, q).  Same result as __fdiv_rn for normal operands
// (checked against IEEE division on 16 M pairs incl. near-tie and all-ones-significand cases,
// CPU test suite), but the reciprocal is loop-invariant, so the serial per-column
// chain is ~12 cycles per division instead of ~40.
__device__ __forceinline__ float div_by(float a, float b, float rb) {
  const float q = __fmul_rn(a, rb);
  const float r = __fmaf_rn(-q, b, a);
  return __fmaf_rn(r, rb, q);
}

// One column step inside a 16-column sub-block; J is a compile-time index so w[] stays in registers.
template <int MODE, int NB, typename CodeT, int J>
struct SubSteps {
  static __device__ __forceinline__ void run(float (&w)[SB], const float* Urow0, float* es, const SolveParams& p,
                                             CodeT* codes, int col0, float s, float rs, float z, __half hs, __half hz,
                                             bool active, int r, int lane) {
    const float* Urow = Urow0 + J * NB;               // row (col0 + J) of the diagonal block, starting at column col0
    const float d = Urow[J];
    const float rd = __frcp_rn(d);                    // independent of w: off the serial chain
    const float x = w[J];
    float q = rintf(div_by(x, s, rs));
    q = __fadd_rn(q, z);
    q = fminf(fmaxf(q, 0.f), p.maxq);
    float deq;
    if (MODE == 0) {
      deq = __fmul_rn(__fsub_rn(q, z), s);
    } else {
      const __half hq = __float2half_rn(q);
      deq = __half2float(__hmul(__hsub(hq, hz), hs));
    }
    const float e = div_by(__fsub_rn(x, deq), d, rd);
    es[J * ROWS + lane] = e;
    if (active) {
      const size_t o = (size_t)(p.c0 + col0 + J) * p.R + r;
      if (MODE == 1) p.Wt[o] = x;                     // the reference leaves the updated weight in place
      if (p.quant_t) p.quant_t[o] = deq;
      if (codes) codes[o] = static_cast<CodeT>(q);
      p.err_t[o] = e;
    }
#pragma unroll
    for (int k4 = (J + 1) / 4; k4 < SB / 4; ++k4) {
      const float4 u = *reinterpret_cast<const float4*>(Urow + k4 * 4);
      const float uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int k = k4 * 4 + t;
        if (k > J) {
          if (MODE == 0) w[k] = __fsub_rn(w[k], __fmul_rn(e, uu[t]));
          else w[k] = __fmaf_rn(-e, uu[t], w[k]);
        }
      }
    }
    SubSteps<MODE, NB, CodeT, J + 1>::run(w, Urow0, es, p, codes, col0, s, rs, z, hs, hz, active, r, lane);
  }
};
template <int MODE, int NB, typename CodeT>
struct SubSteps<MODE, NB, CodeT, SB> {
  static __device__ __forceinline__ void run(float (&)[SB], const float*, float*, const SolveParams&, CodeT*, int,
                                             float, float, float, __half, __half, bool, int, int) {}
};

// CTA = 4 warps x 32 rows.  Shared memory: Ub[NB][NB] (diagonal block of U), ws[NB][32] (the
// block's weights, column-major per row tile: conflict-free), es[16][32] (errors of the current
// sub-block), mm[2][4][32] (min/max partials), one mbarrier.
// Per 16-column sub-block: warp 0 runs the serial recurrence on registers (compile-time column
// indices, ~1 K instructions, reused by the rolled loop over sub-blocks -> instruction-cache
// resident), then all 4 warps apply the rank-16 update to the remaining columns of the block,
// each column seeing its updates in ascending j exactly as the reference's per-column loop does.
template <int MODE, int NB, typename CodeT>
__global__ void __launch_bounds__(SOLVE_WARPS * 32) solve_block_kernel(const SolveParams p) {
  extern __shared__ __align__(16) float smem[];
  float* Ub = smem;
  float* ws = Ub + NB * NB;
  float* es = ws + NB * ROWS;
  float* mm = es + SB * ROWS;
  const uint32_t bar = smem_u32(mm + 2 * SOLVE_WARPS * ROWS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * ROWS + lane;
  const bool active = r < p.R;
  const int R = p.R;

  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    fence_proxy_async();
    uint32_t total = 0;
#pragma unroll 1
    for (int j = 0; j < NB; ++j) total += (NB - (j & ~3)) * 4;
    mbar_expect_tx(bar, total);
#pragma unroll 1
    for (int j = 0; j < NB; ++j) {
      const int cs = j & ~3;                                   // 16-byte aligned start column
      bulk_load_1d(smem_u32(Ub + j * NB + cs), p.U + (size_t)(p.c0 + j) * p.C + p.c0 + cs, (NB - cs) * 4, bar);
    }
  }
  // ---- stage the 32 x NB weight tile (coalesced along rows) and partial min/max
  float mx = -INFINITY, mn = INFINITY;
#pragma unroll 4
  for (int k = warp; k < NB; k += SOLVE_WARPS) {
    const float v = active ? p.Wt[(size_t)(p.c0 + k) * R + r] : 0.f;
    ws[k * ROWS + lane] = v;
    mx = fmaxf(mx, v); mn = fminf(mn, v);
  }
  mm[warp * ROWS + lane] = mx;
  mm[(SOLVE_WARPS + warp) * ROWS + lane] = mn;
  __syncthreads();                                             // also publishes the mbarrier init

  float s = 1.f, z = 0.f;
  if (warp == 0) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 1; i < SOLVE_WARPS; ++i) {
        mx = fmaxf(mx, mm[i * ROWS + lane]);
        mn = fminf(mn, mm[(SOLVE_WARPS + i) * ROWS + lane]);
      }
      const float xmax = fmaxf(mx, 0.f);                       // .relu()           (quant.py:114)
      const float xmin = -fmaxf(-mn, 0.f);                     // -(-min).relu()    (quant.py:115)
      s = __fadd_rn(__fdiv_rn(__fsub_rn(xmax, xmin), p.maxq), 1e-12f);
      z = rintf(__fdiv_rn(-xmin, s));
      if (active) { p.scale[r] = s; p.qzero[r] = z; }
    } else if (active) {
      s = p.scale[r]; z = p.qzero[r];
    }
  }
  const float rs = __frcp_rn(s);
  __half hs = __float2half_rn(0.f), hz = hs;
  if (MODE == 1) { hs = __float2half_rn(s); hz = __float2half_rn(z); }
  CodeT* codes = reinterpret_cast<CodeT*>(p.codes_t);

  mbar_wait(bar, 0);

#pragma unroll 1
  for (int sb = 0; sb < NB / SB; ++sb) {
    const int col0 = sb * SB;
    if (warp == 0) {
      float w[SB];
#pragma unroll
      for (int j = 0; j < SB; ++j) w[j] = ws[(col0 + j) * ROWS + lane];
      SubSteps<MODE, NB, CodeT, 0>::run(w, Ub + col0 * NB + col0, es, p, codes, col0, s, rs, z, hs, hz, active, r, lane);
    }
    __syncthreads();
    if (col0 + SB < NB) {
      float e[SB];
#pragma unroll
      for (int j = 0; j < SB; ++j) e[j] = es[j * ROWS + lane];
#pragma unroll 1
      for (int k = col0 + SB + 4 * warp; k < NB; k += 4 * SOLVE_WARPS) {
        float w0 = ws[(k + 0) * ROWS + lane], w1 = ws[(k + 1) * ROWS + lane];
        float w2 = ws[(k + 2) * ROWS + lane], w3 = ws[(k + 3) * ROWS + lane];
#pragma unroll
        for (int j = 0; j < SB; ++j) {
          const float4 u = *reinterpret_cast<const float4*>(Ub + (col0 + j) * NB + k);
          if (MODE == 0) {
            w0 = __fsub_rn(w0, __fmul_rn(e[j], u.x)); w1 = __fsub_rn(w1, __fmul_rn(e[j], u.y));
            w2 = __fsub_rn(w2, __fmul_rn(e[j], u.z)); w3 = __fsub_rn(w3, __fmul_rn(e[j], u.w));
          } else {
            w0 = __fmaf_rn(-e[j], u.x, w0); w1 = __fmaf_rn(-e[j], u.y, w1);
            w2 = __fmaf_rn(-e[j], u.z, w2); w3 = __fmaf_rn(-e[j], u.w, w3);
          }
        }
        ws[(k + 0) * ROWS + lane] = w0; ws[(k + 1) * ROWS + lane] = w1;
        ws[(k + 2) * ROWS + lane] = w2; ws[(k + 3) * ROWS + lane] = w3;
      }
    }
    __syncthreads();
  }
}

template <int MODE, int NB, typename CodeT>
int launch_block(const SolveParams& p, cudaStream_t st) {
  const int smem = (NB * NB + NB * ROWS + SB * ROWS + 2 * SOLVE_WARPS * ROWS) * 4 + 16;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(solve_block_kernel<MODE, NB, CodeT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) !=
        cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  GQ_LAUNCHED();
  solve_block_kernel<MODE, NB, CodeT><<<(p.R + ROWS - 1) / ROWS, SOLVE_WARPS * 32, smem, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
