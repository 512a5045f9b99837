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
//   blocked right-looking upper Cholesky, NB=128: per step ONE fused launch that factors the diagonal block
//     and solves the panel P = R_kk^-T A[k, k+1:] (chol_panel.cu), then  A22 -= P^T P  on the 3xTF32
//     tensor-core GEMM (upper tiles only)
//   trtri_diag_kernel inverts all diagonal blocks in ONE launch (off the critical path), then the blocked
//   triangular inverse by recursive doubling, entirely as "A^T B" tensor-core GEMMs: with V = R^-1 kept in the
//   upper triangle of A and W = V^T kept in a second buffer,
//        T^T = R12^T W11,    W21 = -V22^T T^T,    V12 = W21^T        (batched over all pairs of a level)
//   reversal of W into U (zeros below the diagonal):  U[i][j] = W[D-1-i][D-1-j].
#include "gptq_b200_internal.h"
#include "ptx.cuh"

namespace gq {
namespace {

// ---------------------------------------------------------------- gather
// A[D-1-i][D-1-j] = scale * Hsym[perm i][perm j] (+ damp on the diagonal), Hp[i][j] the same without the reversal.
// H holds valid data in its upper triangle only (K1 mode 0) and a doubly permuted read of it is a 4-byte random access
// per element, so the gather runs in two coalesced passes:
//   symmetrize_kernel   Hs = full symmetric copy of H (32 x 32 tile transposes through shared memory)
//   gather_rows_kernel  one CTA per output row i: row perm[i] of Hs is staged in shared memory with 16-byte loads, the
//                       column permutation is applied on the shared copy, both outputs are written coalesced.
__global__ void __launch_bounds__(256) symmetrize_kernel(const float* __restrict__ H, float* __restrict__ Hs, int D) {
  __shared__ float tile[32][33];
  const int bj = blockIdx.x, bi = blockIdx.y;
  if (bj < bi) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int yy = ty; yy < 32; yy += 8) {
    const int r = bi * 32 + yy, c = bj * 32 + tx;
    float v = 0.f;
    if (r < D && c < D) {
      v = (c >= r) ? H[(size_t)r * D + c] : H[(size_t)c * D + r];       // (c < r only inside diagonal tiles)
      Hs[(size_t)r * D + c] = v;
    }
    tile[yy][tx] = v;
  }
  if (bi == bj) return;
  __syncthreads();
  for (int yy = ty; yy < 32; yy += 8) {
    const int r = bj * 32 + yy, c = bi * 32 + tx;
    if (r < D && c < D) Hs[(size_t)r * D + c] = tile[tx][yy];
  }
}

__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ Hs, int D, const int64_t* __restrict__ perm,
                                                          float scale, float damp, const float* __restrict__ damp_dev,
                                                          float damp_mult, float* __restrict__ Hp, float* __restrict__ Arev) {
  extern __shared__ __align__(16) float row[];
  if (damp_dev) damp = damp_dev[0] * damp_mult;      // damping computed on the device (no host round trip)
  const int i = blockIdx.x;
  const int pi = perm ? (int)perm[i] : i;
  const float4* src = reinterpret_cast<const float4*>(Hs + (size_t)pi * D);      // D % 4 == 0
  for (int k = threadIdx.x; k < D / 4; k += 256) reinterpret_cast<float4*>(row)[k] = src[k];
  __syncthreads();
  float* hp = Hp ? Hp + (size_t)i * D : nullptr;
  float* ar = Arev ? Arev + (size_t)(D - 1 - i) * D + (D - 1) : nullptr;
  for (int j = threadIdx.x; j < D; j += 256) {
    const int pj = perm ? (int)perm[j] : j;
    float v = row[pj] * scale;
    if (i == j) {
      if (v == 0.f) v = 1.f;     // dead column (gptq_py.py:43-44)
      v += damp;                 // gptq_py.py:58
    }
    if (hp) hp[j] = v;
    if (ar) ar[-j] = v;
  }
}

// element-wise form (one random 4-byte read per element): used when no scratch for the symmetric copy is given
__global__ void gather_kernel(const float* __restrict__ H, int D, const int64_t* __restrict__ perm, float scale,
                              float damp, const float* __restrict__ damp_dev, float damp_mult, float* __restrict__ Hp,
                              float* __restrict__ Arev) {
  if (damp_dev) damp = damp_dev[0] * damp_mult;
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

// Invert every diagonal block R_kk (upper) in place, one CTA per block, all blocks in one launch (runs once after the
// factorisation).  Back substitution T[i][c] = (delta_ic - sum_{i<k<=c} S[i][k] T[k][c]) / S[i][i], rows from the bottom:
// a column only depends on its own later rows, so the 8 lanes that share column c split the dot product by k mod 8 and
// combine with three shuffles; no block-wide barrier inside the loop.  T is kept transposed (Tt[c][k], pitch = 8 mod 32:
// the 4 columns x 8 k-lanes of a warp hit 32 different banks).
constexpr int TR_LANES = 8;
constexpr int TR_THREADS = NB * TR_LANES;     // 1024
constexpr int TR_TPITCH = NB + 8;
__global__ void __launch_bounds__(TR_THREADS) trtri_diag_kernel(float* __restrict__ A, float* __restrict__ Wb, int D) {
  extern __shared__ float sm[];
  float* S = sm;                    // [NB][PITCH] upper factor
  float* Tt = sm + NB * PITCH;      // [NB][TR_TPITCH] transposed inverse: Tt[c][k] = T[k][c]
  const int o = blockIdx.x * NB;
  const int n = min(NB, D - o);
  const int tid = threadIdx.x;
  for (int idx = tid; idx < NB * NB; idx += TR_THREADS) {
    const int r = idx >> 7, c = idx & 127;
    float v = (r == c) ? 1.f : 0.f;
    if (r < n && c < n && c >= r) v = A[(size_t)(o + r) * D + o + c];
    S[r * PITCH + c] = v;
    Tt[r * TR_TPITCH + c] = 0.f;
  }
  __syncthreads();
  const int c = tid / TR_LANES, l = tid % TR_LANES;
  float* tc = Tt + c * TR_TPITCH;
  for (int i = NB - 1; i >= 0; --i) {
    float s0 = 0.f, s1 = 0.f;
    if (i < c) {
      const float* si = S + i * PITCH;
      int k = i + 1 + l;
      for (; k + TR_LANES <= c; k += 2 * TR_LANES) {
        s0 = fmaf(si[k], tc[k], s0);
        s1 = fmaf(si[k + TR_LANES], tc[k + TR_LANES], s1);
      }
      if (k <= c) s0 = fmaf(si[k], tc[k], s0);
    }
    float s = s0 + s1;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (i <= c && l == 0) {
      const float rhs = (i == c) ? 1.f : 0.f;
      tc[i] = (rhs - s) / S[i * PITCH + i];
    }
    __syncwarp();                   // column c only depends on later rows of the same column (same warp)
  }
  __syncthreads();
  for (int idx = tid; idx < NB * NB; idx += TR_THREADS) {
    const int r = idx >> 7, cc = idx & 127;
    if (r < n && cc < n) {
      A[(size_t)(o + r) * D + o + cc] = (cc >= r) ? Tt[cc * TR_TPITCH + r] : 0.f;      // V_kk (upper, zeros below)
      Wb[(size_t)(o + r) * D + o + cc] = (cc <= r) ? Tt[r * TR_TPITCH + cc] : 0.f;     // W_kk = V_kk^T (lower, zeros above)
    }
  }
}

// Zero every 128 x 128 block of A strictly below the block diagonal.  After the factorisation that region still holds the
// input matrix; the triangular products read A-operand windows with TF32_A_UPPER, and on CTA-pair tiles (256 rows, one
// shared k range) the first CTA of a pair reads 128 k-rows below its own diagonal block, which must be zeros.
__global__ void __launch_bounds__(256) zero_block_lower_kernel(float* __restrict__ A, int D) {
  const int bj = blockIdx.x, bi = blockIdx.y;
  if (bi <= bj) return;
  const int r0 = bi * NB, c0 = bj * NB;
  for (int idx = threadIdx.x; idx < NB * (NB / 4); idx += 256) {
    const int r = idx >> 5, c4 = (idx & 31) * 4;
    if (r0 + r < D && c0 + c4 < D) *reinterpret_cast<float4*>(A + (size_t)(r0 + r) * D + c0 + c4) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// V12 = W21^T for every pair of a doubling level: Wb[c0 + j][r0 + i] -> A[r0 + i][c0 + j]
struct TransposeBatch {
  int count;
  struct { int r0, c0, n1, n2; } p[TF32_MAX_BATCH];
};
__global__ void transpose_pairs_kernel(const float* __restrict__ Wb, float* __restrict__ A, int D, const TransposeBatch b) {
  __shared__ float tile[32][33];
  pdl_launch_dependents();
  pdl_wait();
  const int r0 = b.p[blockIdx.z].r0, c0 = b.p[blockIdx.z].c0, n1 = b.p[blockIdx.z].n1, n2 = b.p[blockIdx.z].n2;
  const int i0 = blockIdx.x * 32, j0 = blockIdx.y * 32;      // i along n1 (contiguous in Wb), j along n2 (contiguous in A)
  if (i0 >= n1 || j0 >= n2) return;
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int j = j0 + yy, i = i0 + threadIdx.x;
    tile[yy][threadIdx.x] = (i < n1 && j < n2) ? Wb[(size_t)(c0 + j) * D + r0 + i] : 0.f;
  }
  __syncthreads();
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int i = i0 + yy, j = j0 + threadIdx.x;
    if (i < n1 && j < n2) A[(size_t)(r0 + i) * D + c0 + j] = tile[threadIdx.x][yy];
  }
}

// ---------------------------------------------------------------- reversal
// U[i][j] = W[D-1-i][D-1-j] for i <= j, 0 below the diagonal  (U = J V^T J with W = V^T); 16-byte loads and stores
// (D % 4 == 0: the four reversed source elements of an aligned group of U are an aligned group of W).
__global__ void __launch_bounds__(256) reverse_lower_kernel(const float* __restrict__ Wb, float* __restrict__ U, int D) {
  const int i = blockIdx.y;
  const int j4 = (blockIdx.x * 256 + threadIdx.x) * 4;
  if (j4 >= D) return;
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  if (j4 + 3 >= i) {
    const float4 w = *reinterpret_cast<const float4*>(Wb + (size_t)(D - 1 - i) * D + (D - 4 - j4));
    o.x = (i <= j4) ? w.w : 0.f;
    o.y = (i <= j4 + 1) ? w.z : 0.f;
    o.z = (i <= j4 + 2) ? w.y : 0.f;
    o.w = w.x;
  }
  *reinterpret_cast<float4*>(U + (size_t)i * D + j4) = o;
}

// ---------------------------------------------------------------- packed upper triangle (multi-GPU reduction payload)
// Tile (bi, bj >= bi) of 128 x 128 floats -> packed[t], t = bi*nb - bi*(bi-1)/2 + (bj - bi); edge tiles are zero-padded.
template <bool PACK>
__global__ void __launch_bounds__(256) upper_tiles_kernel(float* __restrict__ H, float* __restrict__ packed, int D, int nb) {
  const int bj = blockIdx.x, bi = blockIdx.y;
  if (bj < bi) return;
  const size_t t = (size_t)bi * nb - (size_t)bi * (bi - 1) / 2 + (bj - bi);
  float* tile = packed + t * (size_t)(NB * NB);
  const int r0 = bi * NB, c0 = bj * NB;
  for (int idx = threadIdx.x; idx < NB * (NB / 4); idx += 256) {
    const int r = idx >> 5, c4 = (idx & 31) * 4;
    const bool in = (r0 + r < D) && (c0 + c4 < D);          // D % 4 == 0
    float* h = H + (size_t)(r0 + r) * D + c0 + c4;
    float4* pk = reinterpret_cast<float4*>(tile + r * NB + c4);
    if (PACK) *pk = in ? *reinterpret_cast<const float4*>(h) : make_float4(0.f, 0.f, 0.f, 0.f);
    else if (in) *reinterpret_cast<float4*>(h) = *pk;
  }
}

}  // namespace

size_t upper_packed_floats(int D) {
  const size_t nb = (size_t)(D + NB - 1) / NB;
  return nb * (nb + 1) / 2 * (size_t)(NB * NB);
}

// pack = 1: H's upper 128 x 128 tiles -> packed; pack = 0: packed -> the same tiles of H (everything else untouched)
int upper_tiles(float* H, float* packed, int D, int pack, cudaStream_t st) {
  if (D <= 0 || (D & 3)) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(H) | reinterpret_cast<uintptr_t>(packed)) & 15) return GQ_ERR_ALIGN;
  const int nb = (D + NB - 1) / NB;
  dim3 grid(nb, nb);
  GQ_LAUNCHED();
  if (pack) upper_tiles_kernel<true><<<grid, 256, 0, st>>>(H, packed, D, nb);
  else upper_tiles_kernel<false><<<grid, 256, 0, st>>>(H, packed, D, nb);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

// Hs: optional (D x D) scratch for the symmetric copy (must not alias Hp / Arev); NULL selects the element-wise gather.
// damp_dev: optional device scalar; when given the damping is damp_dev[0] * damp_mult and `damp` is ignored.
int hessian_gather(const float* H, int D, const int64_t* perm, float scale, float damp, const float* damp_dev, float damp_mult,
                   float* Hp, float* Arev, float* Hs, cudaStream_t st) {
  const int row_smem = D * (int)sizeof(float);
  if (Hs && (D & 3) == 0 && row_smem <= 200 * 1024) {
    static PerDeviceInt smem_set;   // function attributes are per device
    if (row_smem > 48 * 1024 && smem_set.get() < row_smem) {
      if (cudaFuncSetAttribute(gather_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, row_smem) != cudaSuccess)
        return GQ_ERR_CUDA;
      smem_set.set(row_smem);
    }
    const int nt = (D + 31) / 32;
    GQ_LAUNCHED();
    symmetrize_kernel<<<dim3(nt, nt), 256, 0, st>>>(H, Hs, D);
    GQ_LAUNCHED();
    gather_rows_kernel<<<D, 256, row_smem, st>>>(Hs, D, perm, scale, damp, damp_dev, damp_mult, Hp, Arev);
    return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
  }
  dim3 block(32, 8), grid((D + 31) / 32, (D + 7) / 8);
  GQ_LAUNCHED();
  gather_kernel<<<grid, block, 0, st>>>(H, D, perm, scale, damp, damp_dev, damp_mult, Hp, Arev);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

// A: (D x D) work buffer holding the (reversed) SPD matrix, overwritten.  U: (D x D) output.
// Wb: (D x D) scratch for W = V^T.  tmp: >= D*D/2 + 32*D floats.  info_dev: device int, set to 1 on a bad pivot.
int chol_inverse_upper(float* A, int D, float* U, float* Wb, float* tmp, int* info_dev, cudaStream_t st) {
  if (D % 16 != 0) return GQ_ERR_SHAPE;
  static PerDeviceFlag attr_set;   // function attributes are per device
  const int trtri_smem = (NB * PITCH + NB * TR_TPITCH) * (int)sizeof(float);
  if (!attr_set) {
    if (cudaFuncSetAttribute(trtri_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, trtri_smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int nb = (D + NB - 1) / NB;
  int rc;
  Tf32Maps maps;
  if ((rc = gemm_tf32x3_maps(&maps, A, D, D, D, A, D, D, D, A, D, D, D)) != GQ_OK) return rc;
  // ---- blocked right-looking upper Cholesky, trailing update in two levels: inside a super-block of `super` steps
  // only the super-block's own later row blocks are brought up to date after every step (the next panels need
  // them); everything behind the super-block gets ONE rank-(super x 128) update when it closes
  const int super = g_use_tf32x3 ? g_solve_super : 1;
  int sb_o = 0, sb_cnt = 0;
  for (int k = 0; k < nb; ++k) {
    const int o = k * NB;
    const int n = min(NB, D - o);
    const int rest = D - o - n;
    // diagonal block factor + panel solve P = R_kk^-T * A[k, k+1:], one launch (chol_panel.cu)
    if ((rc = chol_factor_panel(A, D, o, n, rest, info_dev, st)) != GQ_OK) return rc;
    if (rest <= 0) break;
    const int e = o + n;
    ++sb_cnt;
    if (g_use_tf32x3) {
      if (sb_cnt == super) {
        // A[e:, e:] -= P^T P with P = A[sb_o:e, e:]  (upper tiles only)
        if ((rc = gemm_tf32x3(maps, rest, rest, e - sb_o, sb_o, e, sb_o, e, e, e, -1.f, 1, st)) != GQ_OK) return rc;
        sb_o = e; sb_cnt = 0;
      } else {
        // rows [e, se) of the trailing matrix only:  A[e:se, e:] -= A[o:e, e:se]^T A[o:e, e:]
        const int se = min(sb_o + super * NB, D);
        if ((rc = gemm_tf32x3(maps, se - e, rest, n, o, e, o, e, e, e, -1.f, 1, st)) != GQ_OK) return rc;
      }
    } else {
      float* panel = A + (size_t)o * D + o + n;
      GemmProblem p2{panel, panel, A + (size_t)(o + n) * D + o + n, rest, rest, n, D, D, D};
      if ((rc = sgemm(p2, -1.f, 1.f, GEMM_TRANS_A | GEMM_C_UPPER, st)) != GQ_OK) return rc;
    }
  }
  // ---- all diagonal blocks: V_kk = R_kk^-1 (into A) and W_kk = V_kk^T (into the zeroed Wb), one launch
  if (cudaMemsetAsync(Wb, 0, (size_t)D * D * sizeof(float), st) != cudaSuccess) return GQ_ERR_CUDA;
  GQ_LAUNCHED();
  trtri_diag_kernel<<<nb, TR_THREADS, trtri_smem, st>>>(A, Wb, D);
  if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  if (nb > 1) {
    GQ_LAUNCHED();
    zero_block_lower_kernel<<<dim3(nb, nb), 256, 0, st>>>(A, D);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  }
  // ---- V = R^-1 by recursive doubling over diagonal blocks
  for (int w = 1; w < nb; w *= 2) {
    const int n1 = w * NB;
    // this level's T^T blocks live in tmp as one (sum n2) x n1 matrix
    Tf32Maps m1, m2;        // m1: T^T = R12^T W11 (A-buffer, Wb -> tmp);  m2: W21 = -V22^T T^T (A-buffer, tmp -> Wb)
    const int t_rows = D;   // upper bound of sum n2 (+ slack rows for 32-row store boxes)
    if ((rc = gemm_tf32x3_maps(&m1, A, D, D, D, Wb, D, D, D, tmp, t_rows, n1, n1)) != GQ_OK) return rc;
    if ((rc = gemm_tf32x3_maps(&m2, A, D, D, D, tmp, t_rows, n1, n1, Wb, D, D, D)) != GQ_OK) return rc;
    Tf32Problem g1[TF32_MAX_BATCH], g2[TF32_MAX_BATCH];
    TransposeBatch tb;
    int cnt = 0, troff = 0, max_n2 = 0;
    auto flush = [&]() -> int {
      if (cnt == 0) return GQ_OK;
      int r1 = gemm_tf32x3_batch(m1, g1, cnt, 1.f, TF32_STORE | TF32_B_LOWER, st);
      if (r1 != GQ_OK) return r1;
      r1 = gemm_tf32x3_batch(m2, g2, cnt, -1.f, TF32_STORE | TF32_A_UPPER, st);
      if (r1 != GQ_OK) return r1;
      tb.count = cnt;
      dim3 block(32, 8), grid((n1 + 31) / 32, (max_n2 + 31) / 32, cnt);
      GQ_LAUNCHED();
      if (launch_chain(transpose_pairs_kernel, grid, block, 0, st, g_pdl != 0, (const float*)Wb, A, D, tb) != cudaSuccess) return GQ_ERR_CUDA;
      cnt = 0; troff = 0; max_n2 = 0;
      return GQ_OK;
    };
    for (int s = 0; s + w < nb; s += 2 * w) {
      const int mid = s + w;
      const int e = min(s + 2 * w, nb);
      const int r0 = s * NB, c0 = mid * NB;
      const int n2 = min(e * NB, D) - c0;
      // T^T (n2 x n1) = R12^T W11 :  A-operand R12 = A[r0:r0+n1, c0:c0+n2] (K = n1), B = Wb[r0:r0+n1, r0:r0+n1] (lower)
      g1[cnt] = Tf32Problem{n2, n1, n1, c0, r0, r0, r0, 0, troff};
      // W21 (n2 x n1) = -V22^T T^T :  A-operand V22 = A[c0:c0+n2, c0:c0+n2] (upper, K = n2), B = T^T, C = Wb[c0:, r0:]
      g2[cnt] = Tf32Problem{n2, n1, n2, c0, c0, 0, troff, r0, c0};
      tb.p[cnt].r0 = r0; tb.p[cnt].c0 = c0; tb.p[cnt].n1 = n1; tb.p[cnt].n2 = n2;
      max_n2 = max(max_n2, n2);
      troff += (n2 + 31) & ~31;
      if (++cnt == TF32_MAX_BATCH) { if ((rc = flush()) != GQ_OK) return rc; }
    }
    if ((rc = flush()) != GQ_OK) return rc;
  }
  dim3 grid((D + 1023) / 1024, D);
  GQ_LAUNCHED();
  reverse_lower_kernel<<<grid, 256, 0, st>>>(Wb, U, D);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
