#pragma once
// K3 + K4  blocked GPTQ column solver on the transposed (C, R) layout.
//
// Replaces, per block of <=128 columns, the reference's 2*128 per-column launches
// (fused_quantize_adjust_kernel + vv_mul_sub_kernel, package/gptq/gptq_cpp_kernel.cu:21-158,
// driven by quantize_range, gptq_cpp.cpp:33-84) by ONE launch, and restates the
// fp32 PyTorch recurrence (gptq_py.py:158-193) the parity tests are judged against.
//
// A CTA of 8 warps owns 32 weight rows (lane = row; rows are independent, so there is no
// cross-row communication).  The block's weights sit in shared memory as ws[column][lane]
// (conflict-free); the upper-triangular NB x NB diagonal block of U (= chol(H^-1)) is staged with
// cp.async.bulk row copies into a PACKED layout (for each 16-row sub-block only the columns to its
// right: 36.9 KB instead of 64 KB, so 3 CTAs fit on an SM and R = 11008 is a single wave) and is
// read as warp-wide broadcasts.
// Software pipeline over 16-column sub-blocks:
//   warp 0      runs the serial column recurrence of sub-block sb on registers.  The recurrence is
//               straight-line code (no branches, no stores to global memory, reciprocals of the
//               diagonal precomputed, round-to-nearest-even by the 1.5*2^23 trick), ~45 instructions
//               and ~13 dependent fp32 operations per column; codes, errors (and in mode 1 the
//               updated weights) are left in shared memory
//   all warps   apply the 16 errors to the NEXT sub-block's 16 columns (the only thing the chain waits for)
//   warps 1-7   write sub-block sb's codes / dequantised values / errors to global memory (coalesced)
//               and apply the errors to the far columns WHILE warp 0 runs the recurrence of sub-block sb+1
// Every column still receives its updates in ascending j, exactly as the reference's per-column loop
// does, so mode 0 stays bit-exact inside a block.
//
// MODE 0 ("pure", gptq_py.py:169-193): scale/zero computed at the block start from the
//   *updated* weights (quant.py:114-118), fp32 dequant, update = w - (e*u) with two
//   roundings (torch: matmul then subtract).  Bit-exact inside a block.
// MODE 1 ("kernel path", gptq_cpp_kernel.cu:41-73,113-137): caller-supplied static
//   scale/zero, dequant rounded through fp16, update = fma(-e, u, w).
//
// The lazy-batch trailing update  W[b:, :] -= U[a:b, b:]^T E[a:b, :]  runs after each block on the
// 3xTF32 tensor-core GEMM (gemm_tf32x3.cu); the reference uses an fp32 addmm_ with TF32 off.
#include <cuda_fp16.h>
#include "gptq_b200_internal.h"
#include "ptx.cuh"
#include "gptq_solver_params.h"

namespace gq {

constexpr int SB = 16;          // columns per register sub-block
constexpr int ROWS = 32;        // weight rows per CTA (one per lane)
constexpr int SOLVE_WARPS = 8;

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

// rintf (round half to even, torch.round / quant.py:19) for |t| < 2^22 as two dependent adds; beyond that range
// the result is only approximately t, which the clamp to [0, maxq] that always follows maps to the same code.
__device__ __forceinline__ float rint_magic(float t) {
  return __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
}

template <int MODE>
__device__ __forceinline__ float dequant_code(float q, float s, float z, __half hs, __half hz) {
  if (MODE == 0) return __fmul_rn(__fsub_rn(q, z), s);                       // quant.py:30 in fp32
  return __half2float(__hmul(__hsub(__float2half_rn(q), hz), hs));           // gptq_cpp_kernel.cu:57-65
}

// Shared-memory store the compiler may move OTHER shared-memory loads across (no "memory" clobber).  The recurrence below
// stores an error and a code per column; as plain C++ stores they fence every later load of the U rows (the compiler
// cannot prove that es / qs do not alias Upk), which put an exposed LDS latency on each of the 128 dependent column
// steps.  Readers of es / qs / xs sit behind a barrier (a compiler memory barrier), or add their own.
__device__ __forceinline__ void sts_noalias(float* p, float v) {
  asm volatile("st.shared.f32 [%0], %1;" :: "r"(smem_u32(p)), "f"(v));
}

// The dependent column recurrence of one 16-column sub-block, straight-line (fully unrolled, w[] in registers);
// everything it leaves behind goes to shared memory.  Software-pipelined by hand: row J + 1 of the U sub-block (the
// multipliers of column J + 1's error), its diagonal entry and reciprocal are loaded BEFORE column J is computed and
// stored.  Left to the compiler, the loads of a row sit behind the stores of the previous column (it cannot prove that
// es / qs do not alias Upk) and every one of the 128 column steps of a block waits for a shared-memory round trip.
template <int MODE, int J>
struct SubSteps {
  static __device__ __forceinline__ void load_row(float (&u)[SB], float& d, float& rd, const float* Urow0, int ld, const float* rdg) {
    const float* Urow = Urow0 + J * ld;               // packed row (col0 + J) of the diagonal block, starting at column col0
    d = Urow[J];
    rd = rdg[J];                                      // RN(1 / d), precomputed in the prologue
#pragma unroll
    for (int k4 = (J + 1) / 4; k4 < SB / 4; ++k4) {
      const float4 v = *reinterpret_cast<const float4*>(Urow + k4 * 4);
      u[k4 * 4 + 0] = v.x; u[k4 * 4 + 1] = v.y; u[k4 * 4 + 2] = v.z; u[k4 * 4 + 3] = v.w;
    }
  }
  // uc / d / rd: row J, loaded one step ago
  static __device__ __forceinline__ void run(float (&w)[SB], const float (&uc)[SB], float d, float rd, const float* Urow0, int ld,
                                             const float* rdg, float* es, float* qs, float* xs, float s, float rs, float z,
                                             float maxq, __half hs, __half hz, int lane) {
    float un[SB], dn = 0.f, rdn = 0.f;
    if (J + 1 < SB) SubSteps<MODE, (J + 1 < SB ? J + 1 : J)>::load_row(un, dn, rdn, Urow0, ld, rdg);
    const float x = w[J];
    float q = rint_magic(div_by(x, s, rs));
    q = __fadd_rn(q, z);
    q = fminf(fmaxf(q, 0.f), maxq);
    const float deq = dequant_code<MODE>(q, s, z, hs, hz);
    const float e = div_by(__fsub_rn(x, deq), d, rd);
    // the next column first: the chain waits for it
#pragma unroll
    for (int k = J + 1; k < SB; ++k) {
      if (MODE == 0) w[k] = __fsub_rn(w[k], __fmul_rn(e, uc[k]));
      else w[k] = __fmaf_rn(-e, uc[k], w[k]);
    }
    es[J * ROWS + lane] = e;
    qs[J * ROWS + lane] = q;
    if (MODE == 1) xs[J * ROWS + lane] = x;           // the reference leaves the updated weight in place
    SubSteps<MODE, J + 1>::run(w, un, dn, rdn, Urow0, ld, rdg, es, qs, xs, s, rs, z, maxq, hs, hz, lane);
  }
  // entry point: loads row 0 itself
  static __device__ __forceinline__ void run(float (&w)[SB], const float* Urow0, int ld, const float* rdg, float* es,
                                             float* qs, float* xs, float s, float rs, float z, float maxq, __half hs,
                                             __half hz, int lane) {
    static_assert(J == 0, "entry point");
    float u0[SB], d0, rd0;
    load_row(u0, d0, rd0, Urow0, ld, rdg);
    run(w, u0, d0, rd0, Urow0, ld, rdg, es, qs, xs, s, rs, z, maxq, hs, hz, lane);
  }
};
template <int MODE>
struct SubSteps<MODE, SB> {
  static __device__ __forceinline__ void run(float (&)[SB], const float (&)[SB], float, float, const float*, int, const float*,
                                             float*, float*, float*, float, float, float, float, __half, __half, int) {}
};

// packed upper-triangular storage of the diagonal block: sub-block sb (rows 16 sb .. 16 sb + 15) keeps columns
// >= 16 sb only, row-major with leading dimension NB - 16 sb
template <int NB> __host__ __device__ constexpr int upk_base(int sb) { return 16 * sb * NB - 128 * sb * (sb - 1); }
template <int NB> constexpr int UPK_FLOATS = upk_base<NB>(NB / SB);

// rank-16 update of 4 consecutive columns k..k+3 (all 32 rows of the CTA) with the errors of sub-block col0
template <int MODE>
__device__ __forceinline__ void update4(float* ws, const float* e, const float* Usb, int ld, int col0, int k, int lane) {
  float w0 = ws[(k + 0) * ROWS + lane], w1 = ws[(k + 1) * ROWS + lane];
  float w2 = ws[(k + 2) * ROWS + lane], w3 = ws[(k + 3) * ROWS + lane];
#pragma unroll
  for (int j = 0; j < SB; ++j) {
    const float4 u = *reinterpret_cast<const float4*>(Usb + j * ld + (k - col0));
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

// Shared memory (floats): Upk (packed diagonal block of U) | ws[NB][32] (the block's weights, column-major per
// row tile: conflict-free) | es, qs, xs [2][16][32] (errors / codes / mode-1 weights of the current and previous
// sub-block) | mm[2][8][32] (min/max partials) | rdg[NB] (reciprocals of diag U) | sz[2][32] (scale, zero per row)
// | one mbarrier.
template <int NB> constexpr int SOLVE_SMEM_FLOATS =
    UPK_FLOATS<NB> + NB * ROWS + 3 * 2 * SB * ROWS + 2 * SOLVE_WARPS * ROWS + NB + 2 * ROWS;

template <int MODE, int NB, typename CodeT>
__global__ void __launch_bounds__(SOLVE_WARPS * 32, 3) solve_block_kernel(const SolveParams p) {
  constexpr int NSB = NB / SB;
  extern __shared__ __align__(16) float smem[];
  float* Upk = smem;
  float* ws = Upk + UPK_FLOATS<NB>;
  float* es = ws + NB * ROWS;
  float* qs = es + 2 * SB * ROWS;
  float* xs = qs + 2 * SB * ROWS;
  float* mm = xs + 2 * SB * ROWS;
  float* rdg = mm + 2 * SOLVE_WARPS * ROWS;
  float* sz = rdg + NB;
  const uint32_t bar = smem_u32(sz + 2 * ROWS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * ROWS + lane;
  const bool active = r < p.R;
  const int R = p.R;

  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    fence_proxy_async();
    mbar_expect_tx(bar, UPK_FLOATS<NB> * 4);
  }
  __syncthreads();                                             // the barrier is initialised and armed
  if (warp == SOLVE_WARPS - 1) {
    // the 128 row copies are issued by 32 lanes in parallel (a single issuing thread costs ~3 us)
    for (int j = lane; j < NB; j += 32) {
      const int sb = j / SB, cs = sb * SB;                     // first column kept for this row
      bulk_load_1d(smem_u32(Upk + upk_base<NB>(sb) + (j - cs) * (NB - cs)),
                   p.U + (size_t)(p.c0 + j) * p.C + p.c0 + cs, (NB - cs) * 4, bar);
    }
  }
  // reciprocals of the diagonal, correctly rounded (needed by div_by), off the serial chain
  if (threadIdx.x < NB) rdg[threadIdx.x] = __frcp_rn(p.U[(size_t)(p.c0 + threadIdx.x) * p.C + p.c0 + threadIdx.x]);
  // ---- stage the 32 x NB weight tile (coalesced along rows) and partial min/max
  float mx = -INFINITY, mn = INFINITY;
  {
    float v[NB / SOLVE_WARPS];
#pragma unroll
    for (int i = 0; i < NB / SOLVE_WARPS; ++i)
      v[i] = active ? p.Wt[(size_t)(p.c0 + warp + i * SOLVE_WARPS) * R + r] : 0.f;
#pragma unroll
    for (int i = 0; i < NB / SOLVE_WARPS; ++i) {
      ws[(warp + i * SOLVE_WARPS) * ROWS + lane] = v[i];
      mx = fmaxf(mx, v[i]); mn = fminf(mn, v[i]);
    }
  }
  mm[warp * ROWS + lane] = mx;
  mm[(SOLVE_WARPS + warp) * ROWS + lane] = mn;
  __syncthreads();

  float s = 1.f, z = 0.f;
  if (warp == 0) {
    if (MODE == 0 && !p.static_params) {
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
    sz[lane] = s; sz[ROWS + lane] = z;
  }
  const float rs = __frcp_rn(s);
  __half hs = __float2half_rn(0.f), hz = hs;
  if (MODE == 1) { hs = __float2half_rn(s); hz = __float2half_rn(z); }
  CodeT* codes = reinterpret_cast<CodeT*>(p.codes_t);
  const float maxq = p.maxq;

  mbar_wait(bar, 0);
  __syncthreads();                                             // sz published (and everybody has seen U)
  if (warp != 0) {
    s = sz[lane]; z = sz[ROWS + lane];
    if (MODE == 1) { hs = __float2half_rn(s); hz = __float2half_rn(z); }
  }

  // global write-out of one finished sub-block, by warps 1..7 (lane = row, coalesced 128-byte rows)
  auto store_subblock = [&](int sb) {
    const float* e_ = es + (sb & 1) * SB * ROWS;
    const float* q_ = qs + (sb & 1) * SB * ROWS;
    const float* x_ = xs + (sb & 1) * SB * ROWS;
    if (!active) return;
    for (int j = warp - 1; j < SB; j += SOLVE_WARPS - 1) {
      const size_t o = (size_t)(p.c0 + sb * SB + j) * R + r;
      const float q = q_[j * ROWS + lane];
      if (MODE == 1) p.Wt[o] = x_[j * ROWS + lane];
      if (p.quant_t) p.quant_t[o] = dequant_code<MODE>(q, s, z, hs, hz);
      if (codes) codes[o] = static_cast<CodeT>(q);
      p.err_t[o] = e_[j * ROWS + lane];
    }
  };

#pragma unroll 1
  for (int sb = 0; sb < NSB; ++sb) {
    const int col0 = sb * SB;
    const int ld = NB - col0;
    const float* Usb = Upk + upk_base<NB>(sb);
    float* esb = es + (sb & 1) * SB * ROWS;
    if (warp == 0) {
      float w[SB];
#pragma unroll
      for (int j = 0; j < SB; ++j) w[j] = ws[(col0 + j) * ROWS + lane];
      SubSteps<MODE, 0>::run(w, Usb, ld, rdg + col0, esb, qs + (sb & 1) * SB * ROWS, xs + (sb & 1) * SB * ROWS, s, rs, z,
                             maxq, hs, hz, lane);
    }
    __syncthreads();                                           // errors / codes of sub-block sb published
    if (sb + 1 == NSB) {
      if (warp > 0) store_subblock(sb);
      break;
    }
    float e[SB];
#pragma unroll
    for (int j = 0; j < SB; ++j) e[j] = esb[j * ROWS + lane];
    // the next sub-block's 16 columns first (4 groups of 4 on warps 0-3): the recurrence waits only for these
    if (warp < SB / 4) update4<MODE>(ws, e, Usb, ld, col0, col0 + SB + 4 * warp, lane);
    __syncthreads();
    // write-out + far columns on warps 1..7 while warp 0 moves on to the next sub-block
    if (warp > 0) {
      store_subblock(sb);
#pragma unroll 1
      for (int k = col0 + 2 * SB + 4 * (warp - 1); k < NB; k += 4 * (SOLVE_WARPS - 1))
        update4<MODE>(ws, e, Usb, ld, col0, k, lane);
    }
  }
}

template <int MODE, int NB, typename CodeT>
int launch_block(const SolveParams& p, cudaStream_t st) {
  const int smem = SOLVE_SMEM_FLOATS<NB> * 4 + 16;
  static PerDeviceFlag attr_set;   // function attributes are per device
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
