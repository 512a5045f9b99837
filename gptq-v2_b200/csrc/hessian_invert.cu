// K2  damp + Cholesky + inverse factor:  U upper with U^T U = H^-1
//
// Replaces HessianHook.invert's three LAPACK calls (gptq_py.py:59-61:
// cholesky_ex(lower) -> cholesky_inverse -> cholesky_ex(upper), 4/3 D^3 flops) by
// 2/3 D^3 flops:  with J the index reversal,  J H J = Rt^T Rt (Rt upper)  =>
// H = S^T S with S = J Rt J lower  =>  H^-1 = (S^-T)^T (S^-T)  and S^-T is upper with
// positive diagonal, so by uniqueness of the Cholesky factor
//        U = J (Rt^-1)^T J,   i.e.  U[i][j] = V[D-1-j][D-1-i],  V = Rt^-1.
// Steps (all fp32, FMA, no TF32):
//   hessian_gather      A[D-1-i][D-1-j] = scale*Hsym[perm i][perm j] (+damp on the diagonal)
//   blocked right-looking upper Cholesky, NB=128: chol_diag_kernel (one CTA, the only serial
//     piece), trsm_panel_kernel  P = R_kk^-T A[k, k+1:]  (forward substitution, 32 columns
//     per CTA), then  A22 -= P^T P  as an "A^T B" GEMM on the upper tiles
//   trtri_diag_kernel inverts all diagonal blocks in ONE launch (off the critical path), then
//   blocked triangular inverse by recursive doubling (batched GEMMs per level)
//   anti-transpose into U (zeros below the diagonal).
#include "gptq_b200_internal.h"

namespace gq {
namespace {

// ---------------------------------------------------------------- gather
__global__ void gather_kernel(const float* __restrict__ H, int D, const int64_t* __restrict__ perm, float scale,
                              float damp, float* __restrict__ Hp, float* __restrict__ Arev) {
  const int j = blockIdx.x * 32 + threadIdx.x;
  const int i = blockIdx.y * 8 + threadIdx.y;
  if (i >= D || j >= D) return;
  const int pi = perm ? (int)perm[i] : i;
  const int pj = perm ? (int)perm[j] : j;
  // H holds valid data in its upper triangle (K1 mode 0); read symmetric
  const float h = (pi <= pj) ? H[(size_t)pi * D + pj] : H[(size_t)pj * D + pi];
  float v = h * scale;
  if (i == j) {
    if (v == 0.f) v = 1.f;     // dead column (gptq_py.py:43-44)
    v += damp;                 // gptq_py.py:58
  }
  if (Hp) Hp[(size_t)i * D + j] = v;
  if (Arev) Arev[(size_t)(D - 1 - i) * D + (D - 1 - j)] = v;
}

// ---------------------------------------------------------------- diagonal block
constexpr int NB = 128;
constexpr int PITCH = NB + 1;

// Factor the n x n (n <= 128) diagonal block at offset o in place:  A_kk = R^T R  (upper R).
// One CTA, 256 threads = 16 x 16, register tiled: thread (ty, tx) owns the 8 x 8 elements
// (ty + 16 i, tx + 16 k) of the block (2-D cyclic, so the shrinking trailing matrix stays
// balanced).  Column step j: every thread applies the rank-1 update of the (already scaled) row j
// to its registers; the half-warp that owns row j+1 updates that row FIRST, takes the pivot with
// one shuffle, and publishes the scaled row j+1 (double-buffered shared memory, one barrier per
// step) while the other warps are still busy with their FMAs.  Row blocks that are finished are
// skipped.  This is the only serial piece of a Cholesky step, so it does nothing else.
__device__ __forceinline__ void chol_publish_row(float (&a)[8][8], int jn, int n, int tx, float* rb, float* gout,
                                                 bool& bad) {
  // caller guarantees: this thread's ty == (jn & 15); the 16 such threads form one half-warp
  const int i1 = jn >> 4;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (i == i1) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = a[i][k];
    }
  float pv = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (k == i1) pv = v[k];                                  // the diagonal element lives where tx == (jn & 15)
  const unsigned half_mask = (jn & 1) ? 0xFFFF0000u : 0x0000FFFFu;    // lane = (ty & 1) * 16 + tx, ty = jn & 15
  const int src_lane = ((jn & 15) & 1) * 16 + (jn & 15);
  const float piv = __shfl_sync(half_mask, pv, src_lane);
  if (jn < n && !(piv > 0.f)) bad = true;                    // also catches NaN
  const float d = sqrtf(piv);
  const float rinv = 1.0f / d;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = tx + 16 * k;
    const float val = (c > jn) ? v[k] * rinv : (c == jn ? d : 0.f);
    rb[c] = val;
    if (jn < n && c >= jn && c < n) gout[c] = val;
  }
}

__global__ void __launch_bounds__(256) chol_diag_kernel(float* __restrict__ A, int D, int o, int n,
                                                        int* __restrict__ info) {
  __shared__ float rowbuf[2][NB];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float a[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int r = ty + 16 * i, c = tx + 16 * k;
      // only the upper triangle is meaningful; pad a partial block with the identity
      a[i][k] = (r < n && c < n) ? ((c >= r) ? A[(size_t)(o + r) * D + o + c] : 0.f) : (r == c ? 1.f : 0.f);
    }
  bool bad = false;
  // NB: the half-warp test below must use the same lane formula as chol_publish_row
  if (ty == 0) chol_publish_row(a, 0, n, tx, rowbuf[0], A + (size_t)o * D + o, bad);
  __syncthreads();
#pragma unroll 1
  for (int j = 0; j < NB - 1; ++j) {
    const float* rb = rowbuf[j & 1];
    float sc[8], sr[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = tx + 16 * k;
      sc[k] = (c > j) ? rb[c] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = ty + 16 * i;
      sr[i] = (r > j) ? rb[r] : 0.f;
    }
    const int jn = j + 1;
    const int i1 = jn >> 4;
    const bool pub = (ty == (jn & 15));
    if (pub) {
      // row j+1 first, so that its pivot chain overlaps the other warps' FMAs
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (i == i1) {
#pragma unroll
          for (int k = 0; k < 8; ++k) a[i][k] = fmaf(-sr[i], sc[k], a[i][k]);
        }
      chol_publish_row(a, jn, n, tx, rowbuf[jn & 1], A + (size_t)(o + jn) * D + o, bad);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (16 * i + 15 > j) {                                   // uniform: block of rows not finished yet
        if (!(pub && i == i1)) {
#pragma unroll
          for (int k = 0; k < 8; ++k) a[i][k] = fmaf(-sr[i], sc[k], a[i][k]);
        }
      }
    }
    __syncthreads();
  }
  if (bad) atomicExch(info, 1);
}

// Panel solve  P = R_kk^-T A[k, k+1:]  in place by forward substitution, 32 panel columns per CTA
// (lane = column, coalesced along the panel row).  Same shape as the GPTQ block solver: warp 0
// runs the 16-row triangular recurrence on registers, all 4 warps apply the rank-16 update to
// the remaining rows.
constexpr int TSB = 16;
template <int J>
struct TrsmSteps {
  static __device__ __forceinline__ void run(float (&w)[TSB], const float* Rrow0, float* xs, float* gout, size_t gstride,
                                             bool active, int lane) {
    const float* Rrow = Rrow0 + J * NB;                // row (row0 + J) of R_kk starting at column row0
    const float x = w[J] / Rrow[J];
    xs[J * 32 + lane] = x;
    if (active) gout[(size_t)J * gstride] = x;
#pragma unroll
    for (int k = J + 1; k < TSB; ++k) w[k] = fmaf(-x, Rrow[k], w[k]);
    TrsmSteps<J + 1>::run(w, Rrow0, xs, gout, gstride, active, lane);
  }
};
template <>
struct TrsmSteps<TSB> {
  static __device__ __forceinline__ void run(float (&)[TSB], const float*, float*, float*, size_t, bool, int) {}
};

__global__ void __launch_bounds__(128) trsm_panel_kernel(float* __restrict__ A, int D, int o, int n, int rest) {
  extern __shared__ __align__(16) float sm[];
  float* Rb = sm;                   // [NB][NB] upper factor of the diagonal block
  float* ws = Rb + NB * NB;         // [NB][32]
  float* xs = ws + NB * 32;         // [16][32]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int col = blockIdx.x * 32 + lane;
  const bool active = col < rest;
  for (int idx = tid; idx < n * (n / 4); idx += 128) {
    const int r = idx / (n / 4), c4 = (idx - r * (n / 4)) * 4;
    *reinterpret_cast<float4*>(Rb + r * NB + c4) =
        *reinterpret_cast<const float4*>(A + (size_t)(o + r) * D + o + c4);
  }
  float* panel = A + (size_t)o * D + o + n + col;
  for (int i = warp; i < n; i += 4) ws[i * 32 + lane] = active ? panel[(size_t)i * D] : 0.f;
  __syncthreads();
#pragma unroll 1
  for (int sb = 0; sb < n / TSB; ++sb) {
    const int row0 = sb * TSB;
    if (warp == 0) {
      float w[TSB];
#pragma unroll
      for (int j = 0; j < TSB; ++j) w[j] = ws[(row0 + j) * 32 + lane];
      TrsmSteps<0>::run(w, Rb + row0 * NB + row0, xs, panel + (size_t)row0 * D, (size_t)D, active, lane);
    }
    __syncthreads();
    if (row0 + TSB < n) {
      float x[TSB];
#pragma unroll
      for (int j = 0; j < TSB; ++j) x[j] = xs[j * 32 + lane];
#pragma unroll 1
      for (int i = row0 + TSB + 4 * warp; i < n; i += 16) {
        float w0 = ws[(i + 0) * 32 + lane], w1 = ws[(i + 1) * 32 + lane];
        float w2 = ws[(i + 2) * 32 + lane], w3 = ws[(i + 3) * 32 + lane];
#pragma unroll
        for (int j = 0; j < TSB; ++j) {
          const float4 u = *reinterpret_cast<const float4*>(Rb + (row0 + j) * NB + i);
          w0 = fmaf(-x[j], u.x, w0); w1 = fmaf(-x[j], u.y, w1);
          w2 = fmaf(-x[j], u.z, w2); w3 = fmaf(-x[j], u.w, w3);
        }
        ws[(i + 0) * 32 + lane] = w0; ws[(i + 1) * 32 + lane] = w1;
        ws[(i + 2) * 32 + lane] = w2; ws[(i + 3) * 32 + lane] = w3;
      }
    }
    __syncthreads();
  }
}

// Invert every diagonal block R_kk (upper) in place, one CTA per block, all blocks in one launch
// (off the critical path: runs once after the factorisation).
__global__ void __launch_bounds__(256) trtri_diag_kernel(float* __restrict__ A, int D) {
  extern __shared__ float sm[];
  float* S = sm;                    // [NB][PITCH] upper factor
  float* Tm = sm + NB * PITCH;      // [NB][PITCH] its inverse
  const int o = blockIdx.x * NB;
  const int n = min(NB, D - o);
  const int tid = threadIdx.x;
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int r = idx >> 7, c = idx & 127;
    float v = (r == c) ? 1.f : 0.f;
    if (r < n && c < n && c >= r) v = A[(size_t)(o + r) * D + o + c];
    S[r * PITCH + c] = v;
    Tm[r * PITCH + c] = 0.f;
  }
  __syncthreads();
  // two adjacent lanes share column c and split the dot product by parity of k
  const int c = tid >> 1, par = tid & 1;
  for (int i = NB - 1; i >= 0; --i) {
    float s0 = 0.f, s1 = 0.f;
    if (c < n && i < c) {
      int k = i + 1 + par;
      for (; k + 2 <= c; k += 4) {
        s0 = fmaf(S[i * PITCH + k], Tm[k * PITCH + c], s0);
        s1 = fmaf(S[i * PITCH + k + 2], Tm[(k + 2) * PITCH + c], s1);
      }
      for (; k <= c; k += 2) s0 = fmaf(S[i * PITCH + k], Tm[k * PITCH + c], s0);
    }
    float s = s0 + s1;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (c < n && i <= c && par == 0) {
      const float rhs = (i == c) ? 1.f : 0.f;
      Tm[i * PITCH + c] = (rhs - s) / S[i * PITCH + i];
    }
    __syncwarp();                   // column c only depends on earlier rows of the same column
  }
  __syncthreads();
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int r = idx >> 7, cc = idx & 127;
    if (r < n && cc < n) A[(size_t)(o + r) * D + o + cc] = (cc >= r) ? Tm[r * PITCH + cc] : 0.f;
  }
}

// ---------------------------------------------------------------- anti-transpose
__global__ void antitranspose_kernel(const float* __restrict__ V, float* __restrict__ U, int D) {
  __shared__ float tile[32][33];
  // output tile (bi, bj) <- input tile rows D-1-j, cols D-1-i
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  // read: input row = D-1-(j0+ty'), input col = D-1-(i0+tx')  (coalesced along tx', descending)
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int j = j0 + yy, i = i0 + threadIdx.x;
    float v = 0.f;
    if (i < D && j < D && i <= j) v = V[(size_t)(D - 1 - j) * D + (D - 1 - i)];
    tile[yy][threadIdx.x] = v;          // tile[j - j0][i - i0]
  }
  __syncthreads();
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int i = i0 + yy, j = j0 + threadIdx.x;
    if (i < D && j < D) U[(size_t)i * D + j] = tile[threadIdx.x][yy];
  }
}

}  // namespace

int hessian_gather(const float* H, int D, const int64_t* perm, float scale, float damp, float* Hp, float* Arev,
                   cudaStream_t st) {
  dim3 block(32, 8), grid((D + 31) / 32, (D + 7) / 8);
  GQ_LAUNCHED();
  gather_kernel<<<grid, block, 0, st>>>(H, D, perm, scale, damp, Hp, Arev);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

// A: (D x D) work buffer holding the (reversed) SPD matrix, overwritten.
// U: (D x D) output.  tmp: >= D*D/2 floats.  info_dev: device int, set to 1 on a bad pivot.
int chol_inverse_upper(float* A, int D, float* U, float* tmp, int* info_dev, cudaStream_t st) {
  if (D % 16 != 0) return GQ_ERR_SHAPE;
  static bool attr_set = false;
  const int diag_smem = 0;
  const int trtri_smem = 2 * NB * PITCH * (int)sizeof(float);
  const int trsm_smem = (NB * NB + NB * 32 + TSB * 32) * (int)sizeof(float);
  if (!attr_set) {
    if (cudaFuncSetAttribute(trtri_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, trtri_smem) != cudaSuccess ||
        cudaFuncSetAttribute(trsm_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, trsm_smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int nb = (D + NB - 1) / NB;
  int rc;
  Tf32Maps maps;
  const bool tc = g_use_tf32x3 != 0;
  if (tc && (rc = gemm_tf32x3_maps(&maps, A, D, D, D, A, D, D, D, A, D, D, D)) != GQ_OK) return rc;
  // ---- blocked right-looking upper Cholesky
  for (int k = 0; k < nb; ++k) {
    const int o = k * NB;
    const int n = min(NB, D - o);
    GQ_LAUNCHED();
    chol_diag_kernel<<<1, 256, diag_smem, st>>>(A, D, o, n, info_dev);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
    const int rest = D - o - n;
    if (rest <= 0) break;
    // P = R_kk^-T * A[k, k+1:]  (forward substitution, in place)
    GQ_LAUNCHED();
    trsm_panel_kernel<<<(rest + 31) / 32, 128, trsm_smem, st>>>(A, D, o, n, rest);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
    // A22 -= P^T P  (upper tiles only)
    float* panel = A + (size_t)o * D + o + n;
    float* a22 = A + (size_t)(o + n) * D + o + n;
    if (tc) {
      if ((rc = gemm_tf32x3(maps, rest, rest, n, o, o + n, o, o + n, o + n, o + n, -1.f, 1, st)) != GQ_OK) return rc;
    } else {
      GemmProblem p2{panel, panel, a22, rest, rest, n, D, D, D};
      if ((rc = sgemm(p2, -1.f, 1.f, GEMM_TRANS_A | GEMM_C_UPPER, st)) != GQ_OK) return rc;
    }
  }
  // ---- all diagonal blocks: R_kk -> R_kk^-1, one launch
  GQ_LAUNCHED();
  trtri_diag_kernel<<<nb, 256, trtri_smem, st>>>(A, D);
  if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  // ---- V = R^-1 by recursive doubling over diagonal blocks: V12 = -V11 (R12 V22)
  for (int w = 1; w < nb; w *= 2) {
    GemmBatch b1, b2;
    b1.count = b2.count = 0;
    size_t toff = 0;
    auto flush = [&]() -> int {
      int r1 = sgemm_batched(b1, 1.f, 0.f, GEMM_B_UPPER, st);
      if (r1 != GQ_OK) return r1;
      int r2 = sgemm_batched(b2, -1.f, 0.f, GEMM_A_UPPER, st);
      b1.count = b2.count = 0;
      toff = 0;
      return r2;
    };
    for (int s = 0; s + w < nb; s += 2 * w) {
      const int mid = s + w;
      const int e = min(s + 2 * w, nb);
      const int r0 = s * NB, c0 = mid * NB;
      const int n1 = w * NB;
      const int n2 = min(e * NB, D) - c0;
      float* t = tmp + toff;
      b1.p[b1.count++] = GemmProblem{A + (size_t)r0 * D + c0, A + (size_t)c0 * D + c0, t, n1, n2, n2, D, D, n2};
      b2.p[b2.count++] = GemmProblem{A + (size_t)r0 * D + r0, t, A + (size_t)r0 * D + c0, n1, n2, n1, D, n2, D};
      toff += (size_t)n1 * n2;
      if (b1.count == GEMM_MAX_BATCH) { if ((rc = flush()) != GQ_OK) return rc; }
    }
    if (b1.count) { if ((rc = flush()) != GQ_OK) return rc; }
  }
  dim3 block(32, 8), grid((D + 31) / 32, (D + 31) / 32);
  GQ_LAUNCHED();
  antitranspose_kernel<<<grid, block, 0, st>>>(A, U, D);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
