// K5  int4 group-quantised storage (pack / unpack / transpose+pack), plus the permuting
// transpose that feeds the solver.  (K6, the W4A16 GEMM that consumes this format, is in w4a16_gemv.cu.)
//
// The reference has neither packing nor a quantised GEMM: it keeps one code per uint8
// (gptq_py.py:217) and evaluates through a dense fp16 reconstruction + cuBLAS
// (quant.py:222-237,262-279; llama.py:172-177).  Storage format defined here:
//   qweight_packed (R, C/8) uint32 : column 8k+i of a row in bits [4i, 4i+4) of word k,
//                                    columns in the act-order-permuted order of quant_meta
//   scale (R, G) fp16/bf16, qzero (R, G) uint8   (as in quant_meta)
// 3-bit codes (maxq = 7) use the same nibble container.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <algorithm>
#include "gptq_b200_internal.h"

namespace gq {
namespace {

// ------------------------------------------------------------------ weight prep
// Wt[c][r] = float(W[r][perm[c]])   (gptq_py.py:90-94 + the .t().contiguous() of :128)
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }

// Wt[c][r] = (float) W[r][perm[c]];  optionally also the same element in W's own type (orig_t, kept for the error
// metrics after the solver has consumed Wt) and colsq[c] += sum_r W[r][perm[c]]^2 (colsq zeroed by the caller).
template <typename T>
__global__ void prep_weight_t_kernel(const T* __restrict__ W, const int64_t* __restrict__ perm, float* __restrict__ Wt,
                                     T* __restrict__ orig_t, float* __restrict__ colsq, int R, int C) {
  const int r = blockIdx.x * 32 + threadIdx.x;
  const int c = blockIdx.y * 8 + threadIdx.y;
  float v = 0.f;
  if (r < R && c < C) {
    const int pc = perm ? (int)perm[c] : c;
    const T w = W[(size_t)r * C + pc];
    v = to_f32<T>(w);
    Wt[(size_t)c * R + r] = v;
    if (orig_t) orig_t[(size_t)c * R + r] = w;
  }
  if (colsq) {                                           // a warp = 32 rows of one column
    float sq = v * v;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    if (threadIdx.x == 0 && c < C) atomicAdd(colsq + c, sq);
  }
}

// Scatter form of the same pass, used whenever the inverse permutation is at hand (HessianHook.perm_inv):
//   Wt[pinv[k]][r] = (float) W[r][k]
// The gather form above reads one 2-byte element per 32-byte sector (a warp touches 32 rows of one permuted column);
// here W is read in its own order (128 B per warp and row), transposed through shared memory, and every destination row
// segment Wt[c][r0 .. r0+31] is one full 128-byte line, wherever c = pinv[k] lands.  Tile = 32 rows x 64 source columns.
template <typename T> struct Pair;
template <> struct Pair<__half> { using type = __half2; };
template <> struct Pair<__nv_bfloat16> { using type = __nv_bfloat162; };
template <> struct Pair<float> { using type = float2; };

// Tile of PREP_ROWS weight rows x 64 source columns: W is read in its own order (4- / 8-byte loads along a row), every
// destination column gets PREP_ROWS contiguous floats (512 B) and one atomic for its sum of squares.
constexpr int PREP_ROWS = 128;
template <typename T>
__global__ void __launch_bounds__(256) prep_weight_scatter_kernel(const T* __restrict__ W, const int64_t* __restrict__ pinv,
                                                                  float* __restrict__ Wt, T* __restrict__ orig_t,
                                                                  float* __restrict__ colsq, int R, int C) {
  using T2 = typename Pair<T>::type;
  __shared__ T tile[64][PREP_ROWS + 2];            // [source column][row], pitch + 2: 2-element stores hit distinct banks
  const int r0 = blockIdx.x * PREP_ROWS, k0 = blockIdx.y * 64;
  const int tx = threadIdx.x, ty = threadIdx.y;    // (32, 8)
  const int k = k0 + 2 * tx;
#pragma unroll 4
  for (int yy = ty; yy < PREP_ROWS; yy += 8) {
    const int r = r0 + yy;
    T a = T(0.f), b = T(0.f);
    if (r < R && k + 1 < C) {                      // C is even (checked by the launcher)
      const T2 v = *reinterpret_cast<const T2*>(W + (size_t)r * C + k);
      a = v.x; b = v.y;
    }
    tile[2 * tx][yy] = a;
    tile[2 * tx + 1][yy] = b;
  }
  __syncthreads();
#pragma unroll 2
  for (int kk = ty; kk < 64; kk += 8) {
    const int ks = k0 + kk;                        // warp-uniform
    if (ks >= C) break;
    const int c = pinv ? (int)pinv[ks] : ks;
    float sq = 0.f;
#pragma unroll
    for (int q = 0; q < PREP_ROWS / 32; ++q) {
      const int r = r0 + tx + 32 * q;
      const T w = tile[kk][tx + 32 * q];
      const float v = to_f32<T>(w);
      if (r < R) {
        Wt[(size_t)c * R + r] = v;
        if (orig_t) orig_t[(size_t)c * R + r] = w;
      }
      sq = fmaf(v, v, sq);                         // rows past R hold zeros
    }
    if (colsq) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
      if (tx == 0) atomicAdd(colsq + c, sq);
    }
  }
}

// out[0] += sum err^2,  out[1] += sum (quant - orig)^2   over n elements of the (C, R) solver outputs
template <typename T>
__global__ void solve_metrics_kernel(const float* __restrict__ err_t, const float* __restrict__ quant_t,
                                     const T* __restrict__ orig_t, size_t n, double* __restrict__ out) {
  float e2 = 0.f, d2 = 0.f;
  const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {   // n % 4 == 0
    const float4 e = *reinterpret_cast<const float4*>(err_t + i);
    const float4 q = *reinterpret_cast<const float4*>(quant_t + i);
    const float o0 = to_f32<T>(orig_t[i]), o1 = to_f32<T>(orig_t[i + 1]), o2 = to_f32<T>(orig_t[i + 2]), o3 = to_f32<T>(orig_t[i + 3]);
    e2 += e.x * e.x + e.y * e.y + e.z * e.z + e.w * e.w;
    const float a = q.x - o0, b = q.y - o1, c = q.z - o2, d = q.w - o3;
    d2 += a * a + b * b + c * c + d * d;
  }
  __shared__ float se[32], sd[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { e2 += __shfl_xor_sync(0xffffffffu, e2, o); d2 += __shfl_xor_sync(0xffffffffu, d2, o); }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { se[warp] = e2; sd[warp] = d2; }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    e2 = lane < nw ? se[lane] : 0.f;
    d2 = lane < nw ? sd[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { e2 += __shfl_xor_sync(0xffffffffu, e2, o); d2 += __shfl_xor_sync(0xffffffffu, d2, o); }
    if (lane == 0) { atomicAdd(out, (double)e2); atomicAdd(out + 1, (double)d2); }
  }
}

// ------------------------------------------------------------------ pack / unpack
__global__ void pack_kernel(const uint8_t* __restrict__ codes, uint32_t* __restrict__ packed, size_t n_words) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_words) return;
  const uint2 v = *reinterpret_cast<const uint2*>(codes + i * 8);
  // bytes b0..b7 -> nibbles
  uint32_t w = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    w |= ((v.x >> (8 * k)) & 0xFu) << (4 * k);
    w |= ((v.y >> (8 * k)) & 0xFu) << (16 + 4 * k);
  }
  packed[i] = w;
}

__global__ void unpack_kernel(const uint32_t* __restrict__ packed, uint8_t* __restrict__ codes, size_t n_words) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_words) return;
  const uint32_t w = packed[i];
  uint2 v;
  v.x = v.y = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    v.x |= ((w >> (4 * k)) & 0xFu) << (8 * k);
    v.y |= ((w >> (16 + 4 * k)) & 0xFu) << (8 * k);
  }
  *reinterpret_cast<uint2*>(codes + i * 8) = v;
}

// Dense 16-bit reconstruction from the packed format, for batches large enough to be compute-bound (the W4A16 GEMM
// re-dequantises the weights for every 256 batch columns; past a few hundred rows one dequantisation + a dense GEMM is
// cheaper):  out[n][k] = T( float(q[n][k] - z[g][n]) * float(s[g][n]) ), one rounding, exactly what the reference's
// construct_matrix stores (quant.py:222-237: int16 subtract, multiply in the 16-bit dtype); permuted column order kept
// (the caller permutes x instead, like K6).  One thread per packed word: 4-byte loads and 16-byte stores, both coalesced.
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }

template <typename T>
__global__ void dequant_w4_kernel(const uint32_t* __restrict__ qw, const T* __restrict__ scale_t,
                                  const uint8_t* __restrict__ qzero_t, T* __restrict__ out, int N, int K, int group) {
  const int kw = K >> 3;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)N * kw) return;
  const int n = (int)(idx / kw), w = (int)(idx - (size_t)n * kw);
  const int g = (w << 3) / group;
  const float s = to_f32<T>(scale_t[(size_t)g * N + n]);
  const int z = qzero_t[(size_t)g * N + n];
  const uint32_t word = qw[idx];
  __align__(16) T v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = from_f32<T>((float)((int)((word >> (4 * i)) & 0xFu) - z) * s);
  *reinterpret_cast<uint4*>(out + idx * 8) = *reinterpret_cast<const uint4*>(v);
}

// codes_t (C, R) uint8 -> codes (R, C) uint8 (optional) and packed (R, C/8) (optional)
__global__ void transpose_pack_kernel(const uint8_t* __restrict__ codes_t, uint8_t* __restrict__ codes,
                                      uint32_t* __restrict__ packed, int R, int C) {
  __shared__ uint8_t tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int c = c0 + yy, r = r0 + threadIdx.x;
    tile[yy][threadIdx.x] = (c < C && r < R) ? codes_t[(size_t)c * R + r] : 0;
  }
  __syncthreads();
  for (int yy = threadIdx.y; yy < 32; yy += 8) {
    const int r = r0 + yy, c = c0 + threadIdx.x;
    if (r < R && c < C) {
      if (codes) codes[(size_t)r * C + c] = tile[threadIdx.x][yy];
      if (packed && (threadIdx.x & 7) == 0) {
        uint32_t w = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) w |= (uint32_t)(tile[threadIdx.x + i][yy] & 0xF) << (4 * i);
        packed[(size_t)r * (C / 8) + c / 8] = w;
      }
    }
  }
}

// Result assembly (gptq_py.py:196-199, 216-237) behind the solver, replacing five torch passes.  The rows of the solver's
// (C, R) arrays may belong to several linears that were solved as one (they share U: q / k / v, up / gate): every sum is
// taken per row segment.
struct SegTable { int n; int bound[GQ_MAX_SEGMENTS + 1]; };   // rows [bound[s], bound[s + 1]) belong to segment s

//   scale_out (R, G) = T(scale_gr (G, R)^T), qzero_out (R, G) = uint8(qzero_gr^T)        32 x 32 tiles through shared memory
template <typename T>
__global__ void __launch_bounds__(256) finalize_params_kernel(const float* __restrict__ scale_gr, const float* __restrict__ qzero_gr,
                                                              int R, int G, T* __restrict__ scale_out, uint8_t* __restrict__ qzero_out) {
  __shared__ float ts[32][33], tz[32][33];
  const int g0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int yy = ty; yy < 32; yy += 8) {
    const int g = g0 + yy, r = r0 + tx;
    const bool in = g < G && r < R;
    ts[yy][tx] = in ? scale_gr[(size_t)g * R + r] : 0.f;
    tz[yy][tx] = in ? qzero_gr[(size_t)g * R + r] : 0.f;
  }
  __syncthreads();
  for (int yy = ty; yy < 32; yy += 8) {
    const int r = r0 + yy, g = g0 + tx;
    if (r < R && g < G) {
      scale_out[(size_t)r * G + g] = from_f32<T>(ts[tx][yy]);
      qzero_out[(size_t)r * G + g] = (uint8_t)(int)tz[tx][yy];
    }
  }
}

// stats[s] = [sum err^2, sum (quant - w)^2, sum_j sum_r w[j][r]^2 / U[p_j][p_j]^2, sum T(w^2)] over the rows of segment s, all from ONE
// pass over err_t / quant_t / orig_t (C, R).  Row j of orig_t holds original column p_j = perm[j] of W, and the reference divides
// the UNPERMUTED W by the permuted-order diag(U) (gptq_py.py:198), hence U[p_j][p_j]; the squares of the fourth sum are rounded
// to the weight dtype like (weight.abs() ** 2.) in gptq_py.py:237.  blockIdx.y = segment.
template <typename T>
__global__ void __launch_bounds__(256) result_sums_kernel(const float* __restrict__ err_t, const float* __restrict__ quant_t,
                                                          const T* __restrict__ orig_t, int R, int C,
                                                          const int64_t* __restrict__ perm, const float* __restrict__ U,
                                                          const SegTable seg, double* __restrict__ stats) {
  const int s = blockIdx.y;
  const int lo = seg.bound[s], w4 = (seg.bound[s + 1] - lo) >> 2;
  const size_t items = (size_t)C * w4;
  float e2 = 0.f, d2 = 0.f, ns = 0.f, l2 = 0.f;
  for (size_t it = (size_t)blockIdx.x * 256 + threadIdx.x; it < items; it += (size_t)gridDim.x * 256) {
    const int j = (int)(it / w4);
    const size_t i = (size_t)j * R + lo + 4 * (size_t)(it - (size_t)j * w4);
    const float4 e = *reinterpret_cast<const float4*>(err_t + i);
    const float4 q = *reinterpret_cast<const float4*>(quant_t + i);
    const float o0 = to_f32<T>(orig_t[i]), o1 = to_f32<T>(orig_t[i + 1]), o2 = to_f32<T>(orig_t[i + 2]), o3 = to_f32<T>(orig_t[i + 3]);
    e2 += e.x * e.x + e.y * e.y + e.z * e.z + e.w * e.w;
    const float a = q.x - o0, b = q.y - o1, c = q.z - o2, d = q.w - o3;
    d2 += a * a + b * b + c * c + d * d;
    const float s0 = o0 * o0, s1 = o1 * o1, s2 = o2 * o2, s3 = o3 * o3;
    l2 += (to_f32<T>(from_f32<T>(s0)) + to_f32<T>(from_f32<T>(s1))) + (to_f32<T>(from_f32<T>(s2)) + to_f32<T>(from_f32<T>(s3)));
    if (U != nullptr) {
      const size_t pj = perm ? (size_t)perm[j] : (size_t)j;
      const float dj = U[pj * C + pj];
      ns += ((s0 + s1) + (s2 + s3)) / (dj * dj);
    }
  }
  __shared__ float red[4][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    e2 += __shfl_xor_sync(0xffffffffu, e2, o); d2 += __shfl_xor_sync(0xffffffffu, d2, o);
    ns += __shfl_xor_sync(0xffffffffu, ns, o); l2 += __shfl_xor_sync(0xffffffffu, l2, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = e2; red[1][warp] = d2; red[2][warp] = ns; red[3][warp] = l2; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += (double)red[threadIdx.x][w];
    atomicAdd(stats + 4 * s + threadIdx.x, t);
  }
}

// codes_t (C, R) -> codes (R, C) / packed (R, C/8) for R % 4 == 0 and C % 8 == 0: 128 x 128 tiles, 4 x 4 byte blocks are
// transposed in registers (__byte_perm), so global and shared memory only ever see 4-byte accesses: a lane reads 4 rows of
// codes_t (4 bytes along r each), writes four words [r][4 c] to shared memory, and rows of 128 codes leave as 128-byte
// stores (64 bytes for the packed nibbles).
__global__ void __launch_bounds__(256) transpose_pack4_kernel(const uint8_t* __restrict__ codes_t, uint8_t* __restrict__ codes,
                                                              uint32_t* __restrict__ packed, int R, int C) {
  __shared__ uint32_t T[128][33];                  // [r][c / 4]
  const int c0 = blockIdx.x * 128, r0 = blockIdx.y * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jl = lane & 7, il = lane >> 3;         // 8 r-blocks x 4 c-blocks per warp step
#pragma unroll
  for (int step = warp; step < 32; step += 8) {    // 32 steps: (c-block group of 4) x (r-block group of 8)
    const int i = (step >> 2) * 4 + il;            // c-block 0..31
    const int j = (step & 3) * 8 + jl;             // r-block 0..31
    const int c = c0 + 4 * i, r = r0 + 4 * j;
    uint32_t v0 = 0, v1 = 0, v2 = 0, v3 = 0;
    if (r < R) {                                   // R % 4 == 0: the whole word is in range
      if (c + 0 < C) v0 = *reinterpret_cast<const uint32_t*>(codes_t + (size_t)(c + 0) * R + r);
      if (c + 1 < C) v1 = *reinterpret_cast<const uint32_t*>(codes_t + (size_t)(c + 1) * R + r);
      if (c + 2 < C) v2 = *reinterpret_cast<const uint32_t*>(codes_t + (size_t)(c + 2) * R + r);
      if (c + 3 < C) v3 = *reinterpret_cast<const uint32_t*>(codes_t + (size_t)(c + 3) * R + r);
    }
    // v_k = bytes (r..r+3) of column c + k  ->  w_m = bytes (c..c+3) of row r + m
    const uint32_t t0 = __byte_perm(v0, v1, 0x5140), t1 = __byte_perm(v0, v1, 0x7362);   // [v0.b0 v1.b0 v0.b1 v1.b1], [v0.b2 v1.b2 v0.b3 v1.b3]
    const uint32_t t2 = __byte_perm(v2, v3, 0x5140), t3 = __byte_perm(v2, v3, 0x7362);
    T[4 * j + 0][i] = __byte_perm(t0, t2, 0x5410);
    T[4 * j + 1][i] = __byte_perm(t0, t2, 0x7632);
    T[4 * j + 2][i] = __byte_perm(t1, t3, 0x5410);
    T[4 * j + 3][i] = __byte_perm(t1, t3, 0x7632);
  }
  __syncthreads();
#pragma unroll 4
  for (int rr = warp; rr < 128; rr += 8) {
    const int r = r0 + rr, c = c0 + 4 * lane;
    const uint32_t w = T[rr][lane];
    if (r < R && c < C) {                          // C % 8 == 0: the whole word is in range
      if (codes) *reinterpret_cast<uint32_t*>(codes + (size_t)r * C + c) = w;
    }
    if (packed) {
      const uint32_t nib = (w & 0xFu) | ((w >> 4) & 0xF0u) | ((w >> 8) & 0xF00u) | ((w >> 12) & 0xF000u);
      const uint32_t hi = __shfl_down_sync(0xffffffffu, nib, 1);
      if (!(lane & 1) && r < R && c < C) packed[(size_t)r * (C / 8) + c / 8] = nib | (hi << 16);
    }
  }
}

// Dense reconstruction of one linear from quant_meta as it is (one code per byte, (R, G) scales / zero points, perm_inv):
//   out[r][c] = T(float(q[r][p] - z[r][p / group]) * float(s[r][p / group])),  p = perm_inv[c]      (quant.py:222-237, 262-279)
// One CTA per row: codes, scales and zero points of the row in shared memory, the output written along c.
template <typename T>
__global__ void __launch_bounds__(256) reconstruct_dense_kernel(const uint8_t* __restrict__ q, const T* __restrict__ scale,
                                                                const uint8_t* __restrict__ qzero, const int16_t* __restrict__ perm_inv,
                                                                T* __restrict__ out, int C, int group) {
  extern __shared__ __align__(16) unsigned char rsm[];
  const int G = C / group;
  uint8_t* sq = rsm;                                           // [C]
  float* ss = reinterpret_cast<float*>(rsm + ((C + 15) & ~15));   // [G]
  float* sz = ss + G;                                          // [G]
  const size_t r = blockIdx.x;
  const uint4* q4 = reinterpret_cast<const uint4*>(q + r * C);
  for (int i = threadIdx.x; i < C / 16; i += blockDim.x) reinterpret_cast<uint4*>(sq)[i] = q4[i];
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    ss[g] = to_f32<T>(scale[r * G + g]);
    sz[g] = (float)qzero[r * G + g];
  }
  __syncthreads();
  const bool pow2 = (group & (group - 1)) == 0;
  const int gshift = 31 - __clz(group);
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int p = perm_inv ? (int)perm_inv[c] : c;
    const int g = pow2 ? p >> gshift : p / group;
    out[r * C + c] = from_f32<T>(((float)sq[p] - sz[g]) * ss[g]);
  }
}

}  // namespace

int prep_weight_t(const void* W, int fmt, const int64_t* perm, float* Wt, void* orig_t, float* colsq, int R, int C,
                  cudaStream_t st) {
  dim3 block(32, 8), grid((R + 31) / 32, (C + 7) / 8);
  if (colsq && cudaMemsetAsync(colsq, 0, (size_t)C * sizeof(float), st) != cudaSuccess) return GQ_ERR_CUDA;
  GQ_LAUNCHED();
  if (fmt == 0) prep_weight_t_kernel<__half><<<grid, block, 0, st>>>((const __half*)W, perm, Wt, (__half*)orig_t, colsq, R, C);
  else if (fmt == 1) prep_weight_t_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)W, perm, Wt, (__nv_bfloat16*)orig_t, colsq, R, C);
  else if (fmt == 2) prep_weight_t_kernel<float><<<grid, block, 0, st>>>((const float*)W, perm, Wt, (float*)orig_t, colsq, R, C);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int prep_weight_scatter(const void* W, int fmt, const int64_t* perm_inv, float* Wt, void* orig_t, float* colsq, int R, int C,
                        cudaStream_t st) {
  if (C & 1) return GQ_ERR_SHAPE;
  if (reinterpret_cast<uintptr_t>(W) & 7) return GQ_ERR_ALIGN;
  dim3 block(32, 8), grid((R + PREP_ROWS - 1) / PREP_ROWS, (C + 63) / 64);
  if (colsq && cudaMemsetAsync(colsq, 0, (size_t)C * sizeof(float), st) != cudaSuccess) return GQ_ERR_CUDA;
  GQ_LAUNCHED();
  if (fmt == 0) prep_weight_scatter_kernel<__half><<<grid, block, 0, st>>>((const __half*)W, perm_inv, Wt, (__half*)orig_t, colsq, R, C);
  else if (fmt == 1) prep_weight_scatter_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)W, perm_inv, Wt, (__nv_bfloat16*)orig_t, colsq, R, C);
  else if (fmt == 2) prep_weight_scatter_kernel<float><<<grid, block, 0, st>>>((const float*)W, perm_inv, Wt, (float*)orig_t, colsq, R, C);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int w4a16_dequant(const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, void* out, int N, int K, int group, int fmt,
                  cudaStream_t st) {
  if (N <= 0 || K <= 0) return GQ_OK;
  if ((K & 7) || group < 8 || (group & 7) || (K % group)) return GQ_ERR_SHAPE;   // a packed word never straddles two groups
  if (reinterpret_cast<uintptr_t>(out) & 15) return GQ_ERR_ALIGN;
  const size_t words = (size_t)N * (K / 8);
  const unsigned grid = (unsigned)((words + 255) / 256);
  GQ_LAUNCHED();
  if (fmt == 0) dequant_w4_kernel<__half><<<grid, 256, 0, st>>>(qw, (const __half*)scale_t, qzero_t, (__half*)out, N, K, group);
  else if (fmt == 1) dequant_w4_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(qw, (const __nv_bfloat16*)scale_t, qzero_t, (__nv_bfloat16*)out, N, K, group);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int reconstruct_dense(const uint8_t* q, const void* scale, const uint8_t* qzero, const int16_t* perm_inv, void* out, int R, int C,
                      int group, int fmt, cudaStream_t st) {
  if (R <= 0 || C <= 0) return GQ_OK;
  if ((C & 15) || group < 1 || (C % group) || C > 32767) return GQ_ERR_SHAPE;      // perm_inv is int16 (gptq_py.py:75)
  if (reinterpret_cast<uintptr_t>(q) & 15) return GQ_ERR_ALIGN;
  const size_t smem = ((C + 15) & ~15) + (size_t)(C / group) * 8;
  if (smem > 48 * 1024) return GQ_ERR_SHAPE;
  GQ_LAUNCHED();
  if (fmt == 0) reconstruct_dense_kernel<__half><<<R, 256, smem, st>>>(q, (const __half*)scale, qzero, perm_inv, (__half*)out, C, group);
  else if (fmt == 1) reconstruct_dense_kernel<__nv_bfloat16><<<R, 256, smem, st>>>(q, (const __nv_bfloat16*)scale, qzero, perm_inv, (__nv_bfloat16*)out, C, group);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int solve_metrics(const float* err_t, const float* quant_t, const void* orig_t, int fmt, size_t n, double* out2, cudaStream_t st) {
  if (n % 4 != 0) return GQ_ERR_SHAPE;
  if (cudaMemsetAsync(out2, 0, 2 * sizeof(double), st) != cudaSuccess) return GQ_ERR_CUDA;
  if (n == 0) return GQ_OK;
  const int threads = 256;
  size_t blocks = (n / 4 + threads - 1) / threads;
  if (blocks > 148 * 8) blocks = 148 * 8;
  GQ_LAUNCHED();
  if (fmt == 0) solve_metrics_kernel<__half><<<(unsigned)blocks, threads, 0, st>>>(err_t, quant_t, (const __half*)orig_t, n, out2);
  else if (fmt == 1) solve_metrics_kernel<__nv_bfloat16><<<(unsigned)blocks, threads, 0, st>>>(err_t, quant_t, (const __nv_bfloat16*)orig_t, n, out2);
  else if (fmt == 2) solve_metrics_kernel<float><<<(unsigned)blocks, threads, 0, st>>>(err_t, quant_t, (const float*)orig_t, n, out2);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int pack_int4(const uint8_t* codes, uint32_t* packed, int R, int C, cudaStream_t st) {
  if (C % 8 != 0) return GQ_ERR_SHAPE;
  if (reinterpret_cast<uintptr_t>(codes) & 7) return GQ_ERR_ALIGN;
  const size_t n = (size_t)R * (C / 8);
  if (n == 0) return GQ_OK;
  GQ_LAUNCHED();
  pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(codes, packed, n);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int unpack_int4(const uint32_t* packed, uint8_t* codes, int R, int C, cudaStream_t st) {
  if (C % 8 != 0) return GQ_ERR_SHAPE;
  if (reinterpret_cast<uintptr_t>(codes) & 7) return GQ_ERR_ALIGN;
  const size_t n = (size_t)R * (C / 8);
  if (n == 0) return GQ_OK;
  GQ_LAUNCHED();
  unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(packed, codes, n);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int transpose_pack_int4(const uint8_t* codes_t, uint8_t* codes, uint32_t* packed, int R, int C, cudaStream_t st) {
  if (C % 8 != 0) return GQ_ERR_SHAPE;
  if (R == 0 || C == 0) return GQ_OK;
  GQ_LAUNCHED();
  if ((R & 3) == 0 && (C & 7) == 0 && ((reinterpret_cast<uintptr_t>(codes_t) | reinterpret_cast<uintptr_t>(codes)) & 3) == 0) {
    transpose_pack4_kernel<<<dim3((C + 127) / 128, (R + 127) / 128), 256, 0, st>>>(codes_t, codes, packed, R, C);
    return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
  }
  dim3 block(32, 8), grid((C + 31) / 32, (R + 31) / 32);
  transpose_pack_kernel<<<grid, block, 0, st>>>(codes_t, codes, packed, R, C);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

// Everything gptq_quant needs after the solver, queued with one call: stats (nseg x 4 doubles, overwritten) = per row segment
// [sum err^2, sum (quant - W)^2, gptq_norm sum, sum T(w^2)] and the (R, G) scale / zero-point tables of quant_meta.
int finalize_quant(const float* err_t, const float* quant_t, const void* orig_t, int fmt, int R, int C, const float* scale_gr,
                   const float* qzero_gr, int G, void* scale_out, uint8_t* qzero_out, const int64_t* perm, const float* U,
                   const int* seg_bounds, int nseg, double* stats, cudaStream_t st) {
  if (R <= 0 || C <= 0 || G <= 0 || nseg <= 0 || nseg > GQ_MAX_SEGMENTS) return GQ_ERR_ARG;
  if (fmt < 0 || fmt > 2) return GQ_ERR_DTYPE;
  SegTable seg;
  seg.n = nseg;
  int max_rows = 0;
  for (int s = 0; s <= nseg; ++s) {
    seg.bound[s] = seg_bounds ? seg_bounds[s] : (s == 0 ? 0 : R);
    if ((seg.bound[s] & 3) || (s > 0 && seg.bound[s] <= seg.bound[s - 1])) return GQ_ERR_SHAPE;
    if (s > 0) max_rows = std::max(max_rows, seg.bound[s] - seg.bound[s - 1]);
  }
  if (seg.bound[0] != 0 || seg.bound[nseg] != R) return GQ_ERR_SHAPE;
  if (cudaMemsetAsync(stats, 0, (size_t)nseg * 4 * sizeof(double), st) != cudaSuccess) return GQ_ERR_CUDA;
  const float* e = err_t ? err_t : quant_t;
  const size_t items = (size_t)C * (max_rows / 4);
  const dim3 sgrid((unsigned)std::min<size_t>((items + 255) / 256, 148 * 8), nseg);
  const dim3 pgrid((G + 31) / 32, (R + 31) / 32);
  GQ_LAUNCHED(); GQ_LAUNCHED();
  if (fmt == 0) {
    result_sums_kernel<__half><<<sgrid, 256, 0, st>>>(e, quant_t, (const __half*)orig_t, R, C, perm, U, seg, stats);
    finalize_params_kernel<__half><<<pgrid, 256, 0, st>>>(scale_gr, qzero_gr, R, G, (__half*)scale_out, qzero_out);
  } else if (fmt == 1) {
    result_sums_kernel<__nv_bfloat16><<<sgrid, 256, 0, st>>>(e, quant_t, (const __nv_bfloat16*)orig_t, R, C, perm, U, seg, stats);
    finalize_params_kernel<__nv_bfloat16><<<pgrid, 256, 0, st>>>(scale_gr, qzero_gr, R, G, (__nv_bfloat16*)scale_out, qzero_out);
  } else {
    result_sums_kernel<float><<<sgrid, 256, 0, st>>>(e, quant_t, (const float*)orig_t, R, C, perm, U, seg, stats);
    finalize_params_kernel<float><<<pgrid, 256, 0, st>>>(scale_gr, qzero_gr, R, G, (float*)scale_out, qzero_out);
  }
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
