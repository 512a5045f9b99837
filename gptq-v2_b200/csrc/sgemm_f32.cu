// fp32 SIMT GEMM building block:  C = beta*C + alpha * op(A) * B   (row-major, true fp32 FMA).
//
// Used where the reference runs fp32 cuBLAS/cuSOLVER with TF32 disabled (main.py:12-13):
//   * lazy-batch trailing update  W[b:, :] -= Hinv[a:b, b:]^T * err[a:b, :]   (gptq_cpp.cpp:81-83)
//   * panel solve / trailing update of the blocked Cholesky and the blocked
//     triangular inverse that replace torch.linalg.cholesky_ex + cholesky_inverse
//     (gptq_py.py:59-61).
// 128x128x16 tiles, 256 threads, 8x8 outputs per thread, double-buffered smem.
// Preconditions: M, N, K, lda, ldb, ldc multiples of 4; pointers 16-byte aligned.
#include "gptq_b200_internal.h"

namespace gq {

namespace {
constexpr int BM = 128, BN = 128, BK = 16;

__device__ __forceinline__ float4 ldg4(const float* p) { return *reinterpret_cast<const float4*>(p); }

template <int FLAGS>
__global__ void __launch_bounds__(256) sgemm_kernel(const __grid_constant__ GemmBatch batch, float alpha, float beta) {
  constexpr bool TA = (FLAGS & GEMM_TRANS_A) != 0;
  constexpr bool CU = (FLAGS & GEMM_C_UPPER) != 0;
  constexpr bool AU = (FLAGS & GEMM_A_UPPER) != 0;
  constexpr bool BU = (FLAGS & GEMM_B_UPPER) != 0;
  const GemmProblem& pr = batch.p[blockIdx.z];
  const int M = pr.M, N = pr.N, K = pr.K;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (m0 >= M || n0 >= N) return;
  if (CU && m0 >= n0 + BN) return;

  __shared__ __align__(16) float As[2][BK][BM];
  __shared__ __align__(16) float Bs[2][BK][BN];

  // K range this tile needs (whole BK chunks; element masks handle the boundary chunk)
  int k_lo = 0, k_hi = K;
  if (AU) { if (TA) k_hi = min(k_hi, m0 + BM); else k_lo = max(k_lo, m0); }
  if (BU) k_hi = min(k_hi, n0 + BN);
  k_lo = (k_lo / BK) * BK;

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 ra[2], rb[2];
  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int idx = tid + h * 256;
      if (TA) {
        const int kr = idx >> 5, c4 = (idx & 31) * 4;
        const int k = k0 + kr, m = m0 + c4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < k_hi && m < M) {
          v = ldg4(pr.A + (size_t)k * pr.lda + m);
          if (AU) { if (k > m) v.x = 0.f; if (k > m + 1) v.y = 0.f; if (k > m + 2) v.z = 0.f; if (k > m + 3) v.w = 0.f; }
        }
        ra[h] = v;
      } else {
        const int row = idx >> 2, kq = (idx & 3) * 4;
        const int m = m0 + row, k = k0 + kq;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < M && k < k_hi) {
          v = ldg4(pr.A + (size_t)m * pr.lda + k);
          if (AU) { if (k < m) v.x = 0.f; if (k + 1 < m) v.y = 0.f; if (k + 2 < m) v.z = 0.f; if (k + 3 < m) v.w = 0.f; }
        }
        ra[h] = v;
      }
      {
        const int kr = idx >> 5, c4 = (idx & 31) * 4;
        const int k = k0 + kr, n = n0 + c4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < k_hi && n < N) {
          v = ldg4(pr.B + (size_t)k * pr.ldb + n);
          if (BU) { if (k > n) v.x = 0.f; if (k > n + 1) v.y = 0.f; if (k > n + 2) v.z = 0.f; if (k > n + 3) v.w = 0.f; }
        }
        rb[h] = v;
      }
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int idx = tid + h * 256;
      if (TA) {
        const int kr = idx >> 5, c4 = (idx & 31) * 4;
        *reinterpret_cast<float4*>(&As[buf][kr][c4]) = ra[h];
      } else {
        const int row = idx >> 2, kq = (idx & 3) * 4;
        As[buf][kq + 0][row] = ra[h].x; As[buf][kq + 1][row] = ra[h].y;
        As[buf][kq + 2][row] = ra[h].z; As[buf][kq + 3][row] = ra[h].w;
      }
      const int kr = idx >> 5, c4 = (idx & 31) * 4;
      *reinterpret_cast<float4*>(&Bs[buf][kr][c4]) = rb[h];
    }
  };

  const int nk = (k_hi > k_lo) ? (k_hi - k_lo + BK - 1) / BK : 0;
  if (nk > 0) {
    load_tiles(k_lo);
    store_tiles(0);
    __syncthreads();
    for (int it = 0; it < nk; ++it) {
      const int buf = it & 1;
      if (it + 1 < nk) load_tiles(k_lo + (it + 1) * BK);
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
        const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      if (it + 1 < nk) {
        store_tiles(buf ^ 1);
        __syncthreads();
      }
    }
  }

  // epilogue
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + ((i < 4) ? (ty * 4 + i) : (64 + ty * 4 + i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int n = n0 + jh * 64 + tx * 4;
      if (n >= N) continue;
      float* cp = pr.C + (size_t)m * pr.ldc + n;
      float4 r;
      r.x = alpha * acc[i][jh * 4 + 0]; r.y = alpha * acc[i][jh * 4 + 1];
      r.z = alpha * acc[i][jh * 4 + 2]; r.w = alpha * acc[i][jh * 4 + 3];
      if (beta != 0.f) {
        const float4 c = *reinterpret_cast<const float4*>(cp);
        r.x = fmaf(beta, c.x, r.x); r.y = fmaf(beta, c.y, r.y);
        r.z = fmaf(beta, c.z, r.z); r.w = fmaf(beta, c.w, r.w);
      }
      *reinterpret_cast<float4*>(cp) = r;
    }
  }
}

template <int FLAGS>
int launch(const GemmBatch& b, float alpha, float beta, cudaStream_t st) {
  int maxM = 0, maxN = 0;
  for (int i = 0; i < b.count; ++i) { maxM = max(maxM, b.p[i].M); maxN = max(maxN, b.p[i].N); }
  if (maxM <= 0 || maxN <= 0) return GQ_OK;
  dim3 grid((maxN + BN - 1) / BN, (maxM + BM - 1) / BM, b.count);
  GQ_LAUNCHED();
  sgemm_kernel<FLAGS><<<grid, 256, 0, st>>>(b, alpha, beta);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}
}  // namespace

int sgemm_batched(const GemmBatch& b, float alpha, float beta, int flags, cudaStream_t st) {
  if (b.count <= 0) return GQ_OK;
  if (b.count > GEMM_MAX_BATCH) return GQ_ERR_ARG;
  for (int i = 0; i < b.count; ++i) {
    const GemmProblem& p = b.p[i];
    if ((p.M | p.N | p.K | p.lda | p.ldb | p.ldc) & 3) return GQ_ERR_SHAPE;
    if ((reinterpret_cast<uintptr_t>(p.A) | reinterpret_cast<uintptr_t>(p.B) | reinterpret_cast<uintptr_t>(p.C)) & 15)
      return GQ_ERR_ALIGN;
  }
  switch (flags) {
    case GEMM_TRANS_A: return launch<GEMM_TRANS_A>(b, alpha, beta, st);
    case GEMM_TRANS_A | GEMM_C_UPPER: return launch<GEMM_TRANS_A | GEMM_C_UPPER>(b, alpha, beta, st);
    case GEMM_TRANS_A | GEMM_A_UPPER: return launch<GEMM_TRANS_A | GEMM_A_UPPER>(b, alpha, beta, st);
    case GEMM_B_UPPER: return launch<GEMM_B_UPPER>(b, alpha, beta, st);
    case GEMM_A_UPPER: return launch<GEMM_A_UPPER>(b, alpha, beta, st);
    case 0: return launch<0>(b, alpha, beta, st);
    default: return GQ_ERR_ARG;
  }
}

int sgemm(const GemmProblem& pr, float alpha, float beta, int flags, cudaStream_t st) {
  GemmBatch b;
  b.p[0] = pr;
  b.count = 1;
  return sgemm_batched(b, alpha, beta, flags, st);
}

}  // namespace gq
