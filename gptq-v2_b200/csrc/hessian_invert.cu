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
//   blocked right-looking upper Cholesky, NB=128; the 128x128 diagonal block is
//     factored AND inverted inside one CTA (chol_diag_kernel), which turns the panel
//     solve into a GEMM:  P = T_kk^T A[k, k+1:],  A22 -= P^T P   (both "A^T B" GEMMs)
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

// Factor the n x n (n <= 128) diagonal block at offset o:  A_kk = R^T R  (upper R),
// then overwrite it with T = R^-1 (upper).  One CTA, 256 threads.
__global__ void __launch_bounds__(256) chol_diag_kernel(float* __restrict__ A, int D, int o, int n,
                                                        int* __restrict__ info) {
  extern __shared__ float sm[];
  float* S = sm;                    // [NB][PITCH] upper factor
  float* Tm = sm + NB * PITCH;      // [NB][PITCH] its inverse
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int r = idx >> 7, c = idx & 127;
    float v = (r == c) ? 1.f : 0.f;
    if (r < n && c < n && c >= r) v = A[(size_t)(o + r) * D + o + c];
    S[r * PITCH + c] = v;
    Tm[r * PITCH + c] = 0.f;
  }
  __syncthreads();
  bool bad = false;
  for (int j = 0; j < n; ++j) {
    const float piv = S[j * PITCH + j];
    if (!(piv > 0.f)) bad = true;           // also catches NaN
    const float d = sqrtf(piv);
    __syncthreads();                        // everyone has read the pivot
    for (int c = j + tid; c < n; c += 256) S[j * PITCH + c] = (c == j) ? d : S[j * PITCH + c] / d;
    __syncthreads();
    for (int r = j + 1 + warp; r < n; r += 8) {
      const float f = S[j * PITCH + r];
      for (int c = r + lane; c < n; c += 32) S[r * PITCH + c] = fmaf(-f, S[j * PITCH + c], S[r * PITCH + c]);
    }
    __syncthreads();
  }
  if (bad && tid == 0) atomicExch(info, 1);
  // inverse: two adjacent lanes share column c and split the dot product by parity of k
  {
    const int c = tid >> 1, par = tid & 1;
    for (int i = NB - 1; i >= 0; --i) {
      // all threads run the same trip count (uniform shuffles); inactive ones contribute 0
      float s = 0.f;
      if (c < n && i < c) {
        for (int k = i + 1 + par; k <= c; k += 2) s = fmaf(S[i * PITCH + k], Tm[k * PITCH + c], s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      if (c < n && i <= c && par == 0) {
        const float rhs = (i == c) ? 1.f : 0.f;
        Tm[i * PITCH + c] = (rhs - s) / S[i * PITCH + i];
      }
      __syncwarp();
      // column c only depends on rows of the same column -> no block-level sync needed
    }
  }
  __syncthreads();
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int r = idx >> 7, c = idx & 127;
    if (r < n && c < n) A[(size_t)(o + r) * D + o + c] = (c >= r) ? Tm[r * PITCH + c] : 0.f;
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
  if (D % 4 != 0) return GQ_ERR_SHAPE;
  static bool attr_set = false;
  const int diag_smem = 2 * NB * PITCH * (int)sizeof(float);
  if (!attr_set) {
    if (cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, diag_smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int nb = (D + NB - 1) / NB;
  int rc;
  // ---- blocked upper Cholesky; diagonal blocks end up holding T_kk = R_kk^-1
  for (int k = 0; k < nb; ++k) {
    const int o = k * NB;
    const int n = min(NB, D - o);
    GQ_LAUNCHED();
    chol_diag_kernel<<<1, 256, diag_smem, st>>>(A, D, o, n, info_dev);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
    const int rest = D - o - n;
    if (rest <= 0) break;
    float* panel = A + (size_t)o * D + o + n;
    // P = T_kk^T * A[k, k+1:]   (in place: each CTA owns whole columns, K = n fits one tile pass)
    GemmProblem p1{A + (size_t)o * D + o, panel, panel, n, rest, n, D, D, D};
    if ((rc = sgemm(p1, 1.f, 0.f, GEMM_TRANS_A | GEMM_A_UPPER, st)) != GQ_OK) return rc;
    // A22 -= P^T P  (upper tiles only)
    float* a22 = A + (size_t)(o + n) * D + o + n;
    GemmProblem p2{panel, panel, a22, rest, rest, n, D, D, D};
    if ((rc = sgemm(p2, -1.f, 1.f, GEMM_TRANS_A | GEMM_C_UPPER, st)) != GQ_OK) return rc;
  }
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
