// K5  int4 group-quantised storage (pack / unpack / transpose+pack) and
// K6  W4A16 dequant-fused GEMM for small batch (HBM-bound GEMV-like path),
// plus the permuting transpose that feeds the solver.
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
#include "gptq_b200_internal.h"

namespace gq {
namespace {

// ------------------------------------------------------------------ weight prep
// Wt[c][r] = float(W[r][perm[c]])   (gptq_py.py:90-94 + the .t().contiguous() of :128)
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }

template <typename T>
__global__ void prep_weight_t_kernel(const T* __restrict__ W, const int64_t* __restrict__ perm, float* __restrict__ Wt,
                                     int R, int C) {
  const int r = blockIdx.x * 32 + threadIdx.x;
  const int c = blockIdx.y * 8 + threadIdx.y;
  if (r >= R || c >= C) return;
  const int pc = perm ? (int)perm[c] : c;
  Wt[(size_t)c * R + r] = to_f32<T>(W[(size_t)r * C + pc]);
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

// ------------------------------------------------------------------ W4A16, small M
// y[m][n] = sum_k x[m][perm[k]] * (q[n][k] - z[n][g(k)]) * s[n][g(k)]
constexpr int KC = 2048;              // K columns staged per pass
constexpr int WARPS = 8;
constexpr int RPW = 2;                // output rows per warp per pass

template <typename XT> __device__ __forceinline__ __half to_half(XT v);
template <> __device__ __forceinline__ __half to_half<__half>(__half v) { return v; }
template <> __device__ __forceinline__ __half to_half<__nv_bfloat16>(__nv_bfloat16 v) {
  return __float2half_rn(fminf(fmaxf(__bfloat162float(v), -65504.f), 65504.f));
}
template <typename XT> __device__ __forceinline__ XT from_f32(float v);
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename XT, int MT>
__global__ void __launch_bounds__(WARPS * 32) w4a16_small_m_kernel(
    const XT* __restrict__ x, const uint32_t* __restrict__ qw, const XT* __restrict__ scale,
    const uint8_t* __restrict__ qzero, const int32_t* __restrict__ perm, XT* __restrict__ y, int M, int N, int K,
    int group, int m0) {
  // xs[m][wq][u4] : 16 bytes = the 8 activations matching word wq of uint4 u4, as 4 half2 (c0,c4)(c1,c5)(c2,c6)(c3,c7)
  __shared__ __align__(16) uint4 xs[MT][4][KC / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_base = (blockIdx.x * WARPS + warp) * RPW;
  const int G = K / group;
  const int words_per_row = K / 8;
  float acc[RPW][MT];
#pragma unroll
  for (int i = 0; i < RPW; ++i)
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[i][m] = 0.f;

  for (int k0 = 0; k0 < K; k0 += KC) {
    const int kc = min(KC, K - k0);
    __syncthreads();
    // ---- stage activations (gather through perm, interleave for the nibble extraction order)
    for (int e = threadIdx.x; e < MT * (KC / 8); e += WARPS * 32) {
      const int m = e / (KC / 8);
      const int wi = e % (KC / 8);                 // word index within the chunk
      const int kk = wi * 8;
      __half h[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float v = 0.f;
        __half hv = __float2half_rn(v);
        if (kk + i < kc && m0 + m < M) {
          const int k = k0 + kk + i;
          const int pk = perm ? perm[k] : k;
          hv = to_half<XT>(x[(size_t)(m0 + m) * K + pk]);
        }
        h[i] = hv;
      }
      uint4 o;
      __half2 p0 = __halves2half2(h[0], h[4]), p1 = __halves2half2(h[1], h[5]);
      __half2 p2 = __halves2half2(h[2], h[6]), p3 = __halves2half2(h[3], h[7]);
      o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
      o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
      xs[m][wi & 3][wi >> 2] = o;
    }
    __syncthreads();
    // ---- stream the packed weights
    const int n_u4 = kc / 32;
    for (int u = lane; u < n_u4; u += 32) {
      uint4 wv[RPW];
      float sc[RPW];
      uint32_t zz[RPW];
      const int g = (k0 + u * 32) / group;
#pragma unroll
      for (int i = 0; i < RPW; ++i) {
        const int n = n_base + i;
        if (n < N) {
          wv[i] = *reinterpret_cast<const uint4*>(qw + (size_t)n * words_per_row + (k0 / 8) + u * 4);
          sc[i] = to_f32<XT>(scale[(size_t)n * G + g]);
          zz[i] = qzero[(size_t)n * G + g];
        } else {
          wv[i] = make_uint4(0, 0, 0, 0); sc[i] = 0.f; zz[i] = 0;
        }
      }
#pragma unroll
      for (int i = 0; i < RPW; ++i) {
        // (1024 + z) in both halves: subtracting it from (1024 + q) gives q - z exactly
        const uint32_t zb = 0x64006400u + zz[i] * 0x00010001u;
        const __half2 zh = *reinterpret_cast<const __half2*>(&zb);
        const uint32_t ww[4] = {wv[i].x, wv[i].y, wv[i].z, wv[i].w};
        __half2 a2[MT][2];
#pragma unroll
        for (int m = 0; m < MT; ++m) { a2[m][0] = __float2half2_rn(0.f); a2[m][1] = a2[m][0]; }
#pragma unroll
        for (int wq = 0; wq < 4; ++wq) {
          __half2 qd[4];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const uint32_t b = ((ww[wq] >> (4 * t)) & 0x000F000Fu) | 0x64006400u;
            qd[t] = __hsub2(*reinterpret_cast<const __half2*>(&b), zh);
          }
#pragma unroll
          for (int m = 0; m < MT; ++m) {
            const uint4 xv = xs[m][wq][u];
            const __half2* xp = reinterpret_cast<const __half2*>(&xv);
#pragma unroll
            for (int t = 0; t < 4; ++t) a2[m][wq & 1] = __hfma2(qd[t], xp[t], a2[m][wq & 1]);
          }
        }
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          const float2 f0 = __half22float2(a2[m][0]);
          const float2 f1 = __half22float2(a2[m][1]);
          acc[i][m] = fmaf(sc[i], (f0.x + f0.y) + (f1.x + f1.y), acc[i][m]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < RPW; ++i)
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      float v = acc[i][m];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && n_base + i < N && m0 + m < M) y[(size_t)(m0 + m) * N + n_base + i] = from_f32<XT>(v);
    }
}

template <typename XT>
int launch_w4a16(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm,
                 void* y, int M, int N, int K, int group, cudaStream_t st) {
  const int grid = (N + WARPS * RPW - 1) / (WARPS * RPW);
  const XT* xx = reinterpret_cast<const XT*>(x);
  const XT* ss = reinterpret_cast<const XT*>(scale);
  XT* yy = reinterpret_cast<XT*>(y);
  int m0 = 0;
  while (m0 < M) {
    const int rem = M - m0;
    GQ_LAUNCHED();
    if (rem >= 4) { w4a16_small_m_kernel<XT, 4><<<grid, WARPS * 32, 0, st>>>(xx, qw, ss, qzero, perm, yy, M, N, K, group, m0); m0 += 4; }
    else if (rem >= 2) { w4a16_small_m_kernel<XT, 2><<<grid, WARPS * 32, 0, st>>>(xx, qw, ss, qzero, perm, yy, M, N, K, group, m0); m0 += 2; }
    else { w4a16_small_m_kernel<XT, 1><<<grid, WARPS * 32, 0, st>>>(xx, qw, ss, qzero, perm, yy, M, N, K, group, m0); m0 += 1; }
  }
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

int prep_weight_t(const void* W, int fmt, const int64_t* perm, float* Wt, int R, int C, cudaStream_t st) {
  dim3 block(32, 8), grid((R + 31) / 32, (C + 7) / 8);
  GQ_LAUNCHED();
  if (fmt == 0) prep_weight_t_kernel<__half><<<grid, block, 0, st>>>((const __half*)W, perm, Wt, R, C);
  else if (fmt == 1) prep_weight_t_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)W, perm, Wt, R, C);
  else if (fmt == 2) prep_weight_t_kernel<float><<<grid, block, 0, st>>>((const float*)W, perm, Wt, R, C);
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
  dim3 block(32, 8), grid((C + 31) / 32, (R + 31) / 32);
  GQ_LAUNCHED();
  transpose_pack_kernel<<<grid, block, 0, st>>>(codes_t, codes, packed, R, C);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int w4a16_gemm(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
               int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M <= 0 || N <= 0) return GQ_OK;
  if (group <= 0 || group % 32 != 0 || K % group != 0 || K % 32 != 0) return GQ_ERR_SHAPE;
  if (reinterpret_cast<uintptr_t>(qw) & 15) return GQ_ERR_ALIGN;
  if (fmt == 0) return launch_w4a16<__half>(x, qw, scale, qzero, perm, y, M, N, K, group, st);
  if (fmt == 1) return launch_w4a16<__nv_bfloat16>(x, qw, scale, qzero, perm, y, M, N, K, group, st);
  return GQ_ERR_DTYPE;
}

}  // namespace gq
