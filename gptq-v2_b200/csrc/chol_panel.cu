// K2, critical path of one blocked-Cholesky step:  factor the 128 x 128 diagonal block AND solve the
// panel to its right in ONE launch.
//
//   A_kk = R_kk^T R_kk (upper R_kk),   P = R_kk^-T A[k, k+1:]
//
// Every CTA redundantly factors the (identical) diagonal block in shared memory and extends the
// same elimination to its own PC panel columns, i.e. it eliminates the 128 x (128 + PC) block row
// [A_kk | A_k,cols].  The diagonal factorisation is a latency-bound dependent chain that occupies
// one SM for ~8 us whether or not the other SMs repeat it, so repeating it removes a launch, a
// grid-wide dependency and the global round trip of R_kk from every step of the factorisation
// (previously: chol_diag_kernel 74 us + trsm_panel_kernel 25 us per step).
//
// In-CTA algorithm, 16-column sub-blocks (8 per block), 256 threads:
//   A  warp 0 factors the 16 x 16 diagonal sub-block on registers: lane k owns column k, every lane
//      tracks a private copy of the diagonal, so a step costs one shuffle + rsqrt (Newton-refined)
//      instead of a block-wide barrier per column (~180 clk per column, issue-bound on one warp).
//   B  one thread per remaining column: 16-row forward substitution on registers.
//   C  rank-16 update of the remaining rows (4-row groups per warp, lanes along columns), with
//      look-ahead: warps 0-3 update the next sub-block's 16 rows first, then warp 0 runs A of the
//      next sub-block while the other warps finish the update.
// fp32 FMA throughout (the reference runs LAPACK potrf in fp32; TF32 is off, main.py:12-13).
#include <utility>
#include "gptq_b200_internal.h"
#include "ptx.cuh"

namespace gq {
namespace {

constexpr int NB = 128;
constexpr int SBK = 16;

// compile-time loop: f(std::integral_constant<int, 0>{}) ... f(integral_constant<int, N - 1>{}).  `#pragma unroll` is only a
// request: the unroller gave up on the 16 x 16 triangle of factor_sub (its budget is shared with the rest of the kernel) and
// emulated the register arrays with predicated moves; with constant indices there is nothing left to decide.
template <typename F, int... Is>
__device__ __forceinline__ void static_for_impl(F&& f, std::integer_sequence<int, Is...>) {
  (f(std::integral_constant<int, Is>{}), ...);
}
template <int N, typename F>
__device__ __forceinline__ void static_for(F&& f) { static_for_impl(f, std::make_integer_sequence<int, N>{}); }

// 1/sqrt(p) to ~1 ulp: MUFU.RSQ + one Newton step (off the IEEE sqrt/div subroutines: this sits on
// the serial chain of the factorisation).
__device__ __forceinline__ float rsqrt_nr(float p) {
  const float y0 = rsqrtf(p);
  const float e = fmaf(-p * y0, y0, 1.0f);
  return fmaf(0.5f * y0, e, y0);
}

template <int PC>
__global__ void __launch_bounds__(256, (PC <= 64) ? 2 : 1)
chol_panel_kernel(float* __restrict__ A, int D, int o, int n, int rest, int* __restrict__ info, long long* __restrict__ dbg) {
  constexpr int W = NB + PC;                 // row pitch of the shared block row
  extern __shared__ __align__(16) float sm[];
  float* S = sm;                             // [NB][W]
  float* rinv = sm + NB * W;                 // [NB] reciprocals of the diagonal of R_kk
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pc0 = blockIdx.x * PC;           // first panel column of this CTA (relative to o + n)
  // profiling hook (gq_set_solve_debug, first step of a factorisation only): clock64 stamps of warps 0 and 1 of CTA 0
  int dbg_i = 0;
  auto stamp = [&]() {
    if (dbg != nullptr && blockIdx.x == 0 && lane == 0 && warp < 2) dbg[64 * warp + dbg_i] = clock64();
    ++dbg_i;
  };
  stamp();
  const int pcn = max(0, min(PC, rest - pc0));

  pdl_launch_dependents();
  pdl_wait();
  // ---- stage [A_kk (upper) | panel columns]; a partial block is padded with the identity.  All of a thread's loads
  // are issued before its first store: one global round trip for the whole block row instead of one per float4.
  {
    constexpr int ND = NB * (NB / 4) / 256;          // float4 per thread, diagonal block (16)
    float4 v[ND];
#pragma unroll
    for (int k = 0; k < ND; ++k) {
      const int idx = tid + k * 256;
      const int r = idx / (NB / 4), c4 = (idx - r * (NB / 4)) * 4;
      v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < n && c4 < n) v[k] = *reinterpret_cast<const float4*>(A + (size_t)(o + r) * D + o + c4);   // n % 4 == 0
      if (r >= n) { v[k].x = (c4 == r) ? 1.f : 0.f; v[k].y = (c4 + 1 == r) ? 1.f : 0.f; v[k].z = (c4 + 2 == r) ? 1.f : 0.f; v[k].w = (c4 + 3 == r) ? 1.f : 0.f; }
    }
    if constexpr (PC > 0) {
      constexpr int NP = NB * (PC / 4) / 256;        // float4 per thread, panel (8 for PC = 64)
      float4 pv[NP];
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        const int idx = tid + k * 256;
        const int r = idx / (PC / 4), c4 = (idx - r * (PC / 4)) * 4;
        pv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < n && c4 < pcn) pv[k] = *reinterpret_cast<const float4*>(A + (size_t)(o + r) * D + o + n + pc0 + c4);   // rest % 4 == 0
      }
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        const int idx = tid + k * 256;
        const int r = idx / (PC / 4), c4 = (idx - r * (PC / 4)) * 4;
        *reinterpret_cast<float4*>(S + r * W + NB + c4) = pv[k];
      }
    }
#pragma unroll
    for (int k = 0; k < ND; ++k) {
      const int idx = tid + k * 256;
      const int r = idx / (NB / 4), c4 = (idx - r * (NB / 4)) * 4;
      *reinterpret_cast<float4*>(S + r * W + c4) = v[k];
    }
  }
  __syncthreads();
  stamp();                                   // [1] block row staged

  bool bad = false;
  // ---- A: 16 x 16 diagonal sub-block at (so, so), warp 0 only
  auto factor_sub = [&](int so) __attribute__((always_inline)) {
    const int k = lane & 15;
    // a[i] = M[i][k] (this lane's column); dg[i] = M[i][i] and e[i] = M[i][i + 1] are private copies kept by EVERY lane with
    // the same fmas in the same order as the owning lanes, so the pivot chain  rsqrt -> R[j][j+1] -> M[j+1][j+1]  needs no
    // shuffle: the shuffles (R[j][i] for the rank-1 update) are issued in the shadow of the next rsqrt.
    float a[SBK], dg[SBK], e[SBK];
#pragma unroll
    for (int i = 0; i < SBK; ++i) {
      a[i] = S[(so + i) * W + so + k];
      dg[i] = S[(so + i) * W + so + i];
      e[i] = (i + 1 < SBK) ? S[(so + i) * W + so + i + 1] : 0.f;
    }
    // software pipeline: the reciprocal square root of the NEXT pivot is started as soon as that pivot is final, before
    // this column's shuffles and rank-1 update are issued, so its latency (MUFU + Newton step) is covered by them
    if (!(dg[0] > 0.f)) bad = true;                        // also catches NaN
    float ri = rsqrt_nr(dg[0]);
    static_for<SBK>([&](auto J) {
      constexpr int j = decltype(J)::value;
      const float p = dg[j];
      const float rn = e[j] * ri;                          // R[j][j+1] (== a[j] * ri of lane j + 1)
      float ri_next = 0.f;
      if constexpr (j + 1 < SBK) {
        dg[j + 1] = fmaf(-rn, rn, dg[j + 1]);
        if (!(dg[j + 1] > 0.f)) bad = true;
        ri_next = rsqrt_nr(dg[j + 1]);
      }
      float d = p * ri;
      d = fmaf(fmaf(-d, d, p), 0.5f * ri, d);              // sqrt(p) refined
      const float r = (k > j) ? a[j] * ri : (k == j ? d : 0.f);
      S[(so + j) * W + so + k] = r;                        // (both half-warps store the same value: no divergent region)
      if (k == j) rinv[so + j] = ri;
      float rj[SBK + 1];                                   // R[j][i], i > j
      static_for<SBK>([&](auto I) {
        constexpr int i = decltype(I)::value;
        if constexpr (i == j + 1) rj[i] = rn;
        else if constexpr (i > j + 1) rj[i] = __shfl_sync(0xffffffffu, r, i);
      });
      static_for<SBK>([&](auto I) {
        constexpr int i = decltype(I)::value;
        if constexpr (i > j) a[i] = fmaf(-rj[i], r, a[i]);
        if constexpr (i > j + 1) {
          dg[i] = fmaf(-rj[i], rj[i], dg[i]);
          e[i - 1] = fmaf(-rj[i - 1], rj[i], e[i - 1]);    // M[i-1][i] -= R[j][i-1] R[j][i]
        }
      });
      ri = ri_next;
    });
  };
  // ---- C: rank-16 update of the 4 rows of group g behind sub-block so:  S[i][c] -= sum_j X[j][i] X[j][c]  (c >= i needed)
  auto update_group = [&](int so, int g) {
    const int i0 = so + SBK + 4 * g;
    float4 xi[SBK];
#pragma unroll
    for (int j = 0; j < SBK; ++j) xi[j] = *reinterpret_cast<const float4*>(S + (so + j) * W + i0);
#pragma unroll 1
    for (int c = (i0 & ~31) + lane; c < W; c += 32) {
      float a0 = S[(i0 + 0) * W + c], a1 = S[(i0 + 1) * W + c];
      float a2 = S[(i0 + 2) * W + c], a3 = S[(i0 + 3) * W + c];
#pragma unroll
      for (int j = 0; j < SBK; ++j) {
        const float xc = S[(so + j) * W + c];
        a0 = fmaf(-xi[j].x, xc, a0); a1 = fmaf(-xi[j].y, xc, a1);
        a2 = fmaf(-xi[j].z, xc, a2); a3 = fmaf(-xi[j].w, xc, a3);
      }
      S[(i0 + 0) * W + c] = a0; S[(i0 + 1) * W + c] = a1;
      S[(i0 + 2) * W + c] = a2; S[(i0 + 3) * W + c] = a3;
    }
  };

  // s = -1 is the prologue (factor the first diagonal sub-block): one loop so that factor_sub is inlined at ONE site (two
  // copies of its fully unrolled body exceed the unroller's budget and the second one falls back to a generic loop)
#pragma unroll 1
  for (int s = -1; s < NB / SBK; ++s) {
    const int so = s * SBK;
    int groups = 0;
    if (s >= 0) {
      // ---- B: rows so..so+15 of every later column: forward substitution, one thread per column
      {
        const int c = so + SBK + tid;
        if (c < W) {
          float x[SBK];
#pragma unroll
          for (int j = 0; j < SBK; ++j) x[j] = S[(so + j) * W + c];
#pragma unroll
          for (int j = 0; j < SBK; ++j) {
            x[j] *= rinv[so + j];
#pragma unroll
            for (int i = j + 1; i < SBK; ++i) x[i] = fmaf(-S[(so + j) * W + so + i], x[j], x[i]);
          }
#pragma unroll
          for (int j = 0; j < SBK; ++j) S[(so + j) * W + c] = x[j];
        }
      }
      __syncthreads();
      stamp();                               // B done
      if (s == NB / SBK - 1) break;
      // ---- C with look-ahead: warps 0-3 first bring the NEXT sub-block's 16 rows up to date (one 4-row group each),
      // then warp 0 factors the next diagonal sub-block while the other seven warps update the remaining rows
      groups = (NB - so - SBK) / 4;
      if (warp < 4) {
        update_group(so, warp);
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      stamp();                               // look-ahead rows done (warps 0-3)
    }
    if (warp == 0) {
      factor_sub(so + SBK);
    } else {
#pragma unroll 1
      for (int g = 4 + (warp - 1); g < groups; g += 7) update_group(so, g);
    }
    stamp();                                 // factor (warp 0) / remaining updates (warp 1) done
    __syncthreads();
    stamp();
  }

  // ---- write back: the factor (CTA 0 only) and this CTA's panel columns
  if (blockIdx.x == 0) {
    for (int idx = tid; idx < n * (n / 4); idx += 256) {
      const int r = idx / (n / 4), c4 = (idx - r * (n / 4)) * 4;
      if (c4 + 3 >= r) {                       // only groups that touch the upper triangle; entries below it untouched
        float4 v = *reinterpret_cast<const float4*>(S + r * W + c4);
        float* dst = A + (size_t)(o + r) * D + o + c4;
        if (c4 >= r) *reinterpret_cast<float4*>(dst) = v;
        else {
          if (c4 + 0 >= r) dst[0] = v.x;
          if (c4 + 1 >= r) dst[1] = v.y;
          if (c4 + 2 >= r) dst[2] = v.z;
          if (c4 + 3 >= r) dst[3] = v.w;
        }
      }
    }
    if (bad) atomicExch(info, 1);
  }
  if constexpr (PC > 0) {
    for (int idx = tid; idx < n * (PC / 4); idx += 256) {
      const int r = idx / (PC / 4), c4 = (idx - r * (PC / 4)) * 4;
      if (c4 < pcn)
        *reinterpret_cast<float4*>(A + (size_t)(o + r) * D + o + n + pc0 + c4) =
            *reinterpret_cast<const float4*>(S + r * W + NB + c4);
    }
  }
  stamp();                                   // written back
}

template <int PC>
int launch_cp(float* A, int D, int o, int n, int rest, int* info, cudaStream_t st) {
  const int smem = (NB * (NB + PC) + NB) * (int)sizeof(float);
  static PerDeviceFlag attr_set;   // function attributes are per device
  if (!attr_set) {
    if (cudaFuncSetAttribute(chol_panel_kernel<PC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int grid = PC > 0 ? (rest + PC - 1) / PC : 1;
  GQ_LAUNCHED();
  return launch_chain(chol_panel_kernel<PC>, dim3(grid), dim3(256), (size_t)smem, st, g_pdl != 0, A, D, o, n, rest, info,
                      (o == 0) ? g_solve_dbg : (long long*)nullptr) == cudaSuccess
             ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

// One step of the blocked upper Cholesky of A (D x D, upper triangle meaningful): factor the n x n block at
// (o, o) in place and overwrite A[o:o+n, o+n:] with the solved panel.  rest = D - o - n.
int chol_factor_panel(float* A, int D, int o, int n, int rest, int* info_dev, cudaStream_t st) {
  if ((n & 3) || (rest & 3) || (D & 3)) return GQ_ERR_SHAPE;
  if (rest <= 0) return launch_cp<0>(A, D, o, n, 0, info_dev, st);
  return launch_cp<64>(A, D, o, n, rest, info_dev, st);
}

}  // namespace gq
