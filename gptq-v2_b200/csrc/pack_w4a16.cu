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

}  // namespace gq
