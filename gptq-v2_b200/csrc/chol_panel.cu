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
__global__ void __launch_bounds__(256, 1)
chol_panel_kernel(float* __restrict__ A, int D, int o, int n, int rest, int* __restrict__ info, long long* __restrict__ dbg) {
  constexpr int W = NB + PC;                 // row pitch of the shared block row
  extern __shared__ __align__(16) float sm[];
  float* S = sm;                             // [NB][W]
  float* rinv = sm + NB * W;                 // [NB] reciprocals of the diagonal of R_kk
  int* work_ctr = reinterpret_cast<int*>(sm + NB * W + NB);   // next unclaimed update unit of the current sub-block
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
  // ---- stage [A_kk (upper) | panel columns]; a partial block is padded with the identity.  Warp 0 fetches only what the
  // first diagonal sub-block's factorisation needs (rows 0..15 of A_kk) and starts on it while warps 1-7 stage the rest.
  auto load_diag = [&](int r, int c4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < n && c4 < n) v = *reinterpret_cast<const float4*>(A + (size_t)(o + r) * D + o + c4);   // n % 4 == 0
    if (r >= n) { v.x = (c4 == r) ? 1.f : 0.f; v.y = (c4 + 1 == r) ? 1.f : 0.f; v.z = (c4 + 2 == r) ? 1.f : 0.f; v.w = (c4 + 3 == r) ? 1.f : 0.f; }
    return v;
  };
  if (warp == 0) {
    float4 v[SBK];
#pragma unroll
    for (int k = 0; k < SBK; ++k) v[k] = load_diag(k, 4 * lane);                  // row k, 128 columns across the warp
#pragma unroll
    for (int k = 0; k < SBK; ++k) *reinterpret_cast<float4*>(S + k * W + 4 * lane) = v[k];
    __syncwarp();
  } else {
    const int wt = tid - 32;                                                       // 0..223
    constexpr int DI = (NB - SBK) * (NB / 4);                                      // diagonal block, rows 16..127
    constexpr int UNR = 8;                                                         // loads in flight per thread
    for (int base = 0; base < DI; base += 224 * UNR) {
      float4 v[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int idx = base + wt + k * 224;
        const int r = SBK + idx / (NB / 4), c4 = (idx % (NB / 4)) * 4;
        if (idx < DI) v[k] = load_diag(r, c4);
      }
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int idx = base + wt + k * 224;
        const int r = SBK + idx / (NB / 4), c4 = (idx % (NB / 4)) * 4;
        if (idx < DI) *reinterpret_cast<float4*>(S + r * W + c4) = v[k];
      }
    }
    if constexpr (PC > 0) {
      constexpr int PI = NB * (PC / 4);
      for (int base = 0; base < PI; base += 224 * UNR) {
        float4 v[UNR];
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
          const int idx = base + wt + k * 224;
          const int r = idx / (PC / 4), c4 = (idx % (PC / 4)) * 4;
          v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (idx < PI && r < n && c4 < pcn) v[k] = *reinterpret_cast<const float4*>(A + (size_t)(o + r) * D + o + n + pc0 + c4);   // rest % 4 == 0
        }
#pragma unroll
        for (int k = 0; k < UNR; ++k) {
          const int idx = base + wt + k * 224;
          const int r = idx / (PC / 4), c4 = (idx % (PC / 4)) * 4;
          if (idx < PI) *reinterpret_cast<float4*>(S + r * W + NB + c4) = v[k];
        }
      }
    }
  }
  stamp();                                   // [1] this warp's share of the block row staged (no block-wide barrier here)

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
  // ---- C: rank-16 update of the 8 rows of group g behind sub-block so:  S[i][c] -= sum_j X[j][i] X[j][c]  (c >= i needed),
  // 32-column chunks [first, last) counted from the chunk that holds the group's own diagonal.  The update is bound by
  // shared-memory traffic: per chunk a lane reads 16 values of its column and reads + writes its 8 outputs for 128 fmas
  // (the row factors X[j][i] stay in registers), 1 B/fma against the 1.5 B/fma of 4-row groups.
  auto update_group = [&](int so, int g, int first, int last) {
    const int i0 = so + SBK + 8 * g;
    float4 xa[SBK], xb[SBK];
#pragma unroll
    for (int j = 0; j < SBK; ++j) {
      xa[j] = *reinterpret_cast<const float4*>(S + (so + j) * W + i0);
      xb[j] = *reinterpret_cast<const float4*>(S + (so + j) * W + i0 + 4);
    }
    const int c_begin = (i0 & ~31) + 32 * first + lane;
    const int c_end = min(W, (i0 & ~31) + 32 * last);
#pragma unroll 1
    for (int c = c_begin; c < c_end; c += 32) {
      float a[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) a[r] = S[(i0 + r) * W + c];
#pragma unroll
      for (int j = 0; j < SBK; ++j) {
        const float xc = S[(so + j) * W + c];
        a[0] = fmaf(-xa[j].x, xc, a[0]); a[1] = fmaf(-xa[j].y, xc, a[1]);
        a[2] = fmaf(-xa[j].z, xc, a[2]); a[3] = fmaf(-xa[j].w, xc, a[3]);
        a[4] = fmaf(-xb[j].x, xc, a[4]); a[5] = fmaf(-xb[j].y, xc, a[5]);
        a[6] = fmaf(-xb[j].z, xc, a[6]); a[7] = fmaf(-xb[j].w, xc, a[7]);
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) S[(i0 + r) * W + c] = a[r];
    }
  };

  // s = -1 is the prologue (factor the first diagonal sub-block): one loop so that factor_sub is inlined at ONE site (two
  // copies of its fully unrolled body exceed the unroller's budget and the second one falls back to a generic loop)
#pragma unroll 1
  for (int s = -1; s < NB / SBK; ++s) {
    const int so = s * SBK;
    int groups = 0;
    if (s >= 0) {
      // ---- B: rows so..so+15 of every later column: forward substitution, one thread per column; the 16 x 16 factor is
      // read with 16-byte broadcast loads (36 instead of 120 shared-memory instructions per thread)
      {
        const int c = so + SBK + tid;
        if (c < W) {
          float x[SBK];
#pragma unroll
          for (int j = 0; j < SBK; ++j) x[j] = S[(so + j) * W + c];
          float ri[SBK];
#pragma unroll
          for (int q = 0; q < SBK / 4; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(rinv + so + 4 * q);
            ri[4 * q] = v.x; ri[4 * q + 1] = v.y; ri[4 * q + 2] = v.z; ri[4 * q + 3] = v.w;
          }
          static_for<SBK>([&](auto J) {
            constexpr int j = decltype(J)::value;
            x[j] *= ri[j];
            static_for<SBK / 4>([&](auto Q) {
              constexpr int q = decltype(Q)::value;
              if constexpr (4 * q + 3 > j) {
                const float4 v = *reinterpret_cast<const float4*>(S + (so + j) * W + so + 4 * q);
                if constexpr (4 * q + 0 > j) x[4 * q + 0] = fmaf(-v.x, x[j], x[4 * q + 0]);
                if constexpr (4 * q + 1 > j) x[4 * q + 1] = fmaf(-v.y, x[j], x[4 * q + 1]);
                if constexpr (4 * q + 2 > j) x[4 * q + 2] = fmaf(-v.z, x[j], x[4 * q + 2]);
                if constexpr (4 * q + 3 > j) x[4 * q + 3] = fmaf(-v.w, x[j], x[4 * q + 3]);
              }
            });
          });
#pragma unroll
          for (int j = 0; j < SBK; ++j) S[(so + j) * W + c] = x[j];
        }
        if (tid == 0) *work_ctr = 0;
      }
      __syncthreads();
      stamp();                               // B done
      // rows so..so+15 are final (factor rows and solved panel rows): they leave now, under the updates of the later rows
      // (warps 2-7; warps 0-1 are on the look-ahead's path), instead of in one burst at the end of the kernel
      if (warp >= 2) {
        const int wt = tid - 64;                                                   // 0..191
        if (blockIdx.x == 0) {
          for (int idx = wt; idx < SBK * (NB / 4); idx += 192) {
            const int r = so + idx / (NB / 4), c4 = (idx % (NB / 4)) * 4;
            if (r < n && c4 < n) {
              float4 v = *reinterpret_cast<const float4*>(S + r * W + c4);         // zeros below the diagonal
              v.x = (c4 + 0 >= r) ? v.x : 0.f;
              v.y = (c4 + 1 >= r) ? v.y : 0.f;
              v.z = (c4 + 2 >= r) ? v.z : 0.f;
              v.w = (c4 + 3 >= r) ? v.w : 0.f;
              *reinterpret_cast<float4*>(A + (size_t)(o + r) * D + o + c4) = v;
            }
          }
        }
        if constexpr (PC > 0) {
          for (int idx = wt; idx < SBK * (PC / 4); idx += 192) {
            const int r = so + idx / (PC / 4), c4 = (idx % (PC / 4)) * 4;
            if (r < n && c4 < pcn)
              *reinterpret_cast<float4*>(A + (size_t)(o + r) * D + o + n + pc0 + c4) =
                  *reinterpret_cast<const float4*>(S + r * W + NB + c4);
          }
        }
      }
      if (s == NB / SBK - 1) break;
      // ---- C with look-ahead: warps 0-1 first bring the chunk that holds the NEXT diagonal sub-block up to date (one
      // 8-row group each, one 32-column chunk), then warp 0 factors it while everybody else -- and warp 0 afterwards --
      // claims the remaining update units from a shared counter
      groups = (NB - so - SBK) / 8;
      if (warp < 2) {
        update_group(so, warp, 0, 1);
        asm volatile("bar.sync 1, 64;" ::: "memory");
      }
      stamp();                               // look-ahead chunk done (warps 0-1)
    }
    if (warp == 0) factor_sub(so + SBK);
    if (s >= 0) {
      stamp();                               // factor done (warp 0)
      // units: groups 2.. whole (longest first), then the rest of the two look-ahead groups
      for (;;) {
        int u = 0;
        if (lane == 0) u = atomicAdd(work_ctr, 1);
        u = __shfl_sync(0xffffffffu, u, 0);
        if (u >= groups) break;
        if (u < groups - 2) update_group(so, u + 2, 0, 8);
        else update_group(so, u - (groups - 2), 1, 8);
      }
    }
    stamp();                                 // updates done
    __syncthreads();
    stamp();
  }

  // (every row group was written back when it became final)
  if (blockIdx.x == 0 && bad) atomicExch(info, 1);
  stamp();                                   // written back
}

template <int PC>
int launch_cp(float* A, int D, int o, int n, int rest, int* info, cudaStream_t st) {
  const int smem = (NB * (NB + PC) + NB + 4) * (int)sizeof(float);
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
  // one CTA per SM (the update tiles want > 128 registers): 64 panel columns per CTA while that fits one wave, else 128
  static int n_sms = 0;
  if (n_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sms <= 0) n_sms = 148;
  }
  if ((rest + 63) / 64 > n_sms) return launch_cp<128>(A, D, o, n, rest, info_dev, st);
  return launch_cp<64>(A, D, o, n, rest, info_dev, st);
}

}  // namespace gq
