// Row f1 of SURVEY section 8 (driver costs): the elementwise steps of a Llama decoder block between the quantised linears,
// as the layer-by-layer driver replays them over all calibration tokens (llama.py:121-125, 176-177, 181-189).  With K1-K4
// at 2.3 s per 7B model the plain-PyTorch forward was 5.7 s, a third of it in multi-pass elementwise chains (rms_norm = 7
// passes, rotary = 8, silu * up = 3).  Each is ONE pass here, HBM-bound (algorithmic bytes in the comment of each kernel),
// with the roundings of the PyTorch / HF Llama expressions it replaces, so the Hessians see the same activations.
#include "gptq_b200_internal.h"
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace gq {
namespace {

template <typename T> __device__ __forceinline__ float lf(T v);
template <> __device__ __forceinline__ float lf<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float lf<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T lt(float v);
template <> __device__ __forceinline__ __half lt<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 lt<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// y = w * T(x_f32 * rsqrt(mean(x_f32^2) + eps))      (HF LlamaRMSNorm: normalise in fp32, cast, multiply in the model dtype)
// One warp per row, the row kept in registers (D <= 8192): 2 D bytes read + 2 D written per row.
template <typename T, int VPL>   // VPL: 16-byte vectors per lane
__global__ void __launch_bounds__(256) rms_norm_kernel(const T* x, const T* __restrict__ w, T* y, long long rows,
                                                       int D, float eps) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * D);
  const int nvec = D >> 3;
  uint4 v[VPL];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = lane + 32 * i;
    v[i] = make_uint4(0, 0, 0, 0);
    if (idx < nvec) v[i] = xr[idx];
    const T* e = reinterpret_cast<const T*>(&v[i]);
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float f = lf<T>(e[j]); ss = fmaf(f, f, ss); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float rs = rsqrtf(ss / (float)D + eps);
  uint4* yr = reinterpret_cast<uint4*>(y + row * D);
  const uint4* wr = reinterpret_cast<const uint4*>(w);
#pragma unroll
  for (int i = 0; i < VPL; ++i) {
    const int idx = lane + 32 * i;
    if (idx < nvec) {
      const uint4 wv = wr[idx];
      const T* e = reinterpret_cast<const T*>(&v[i]);
      const T* we = reinterpret_cast<const T*>(&wv);
      __align__(16) T o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = lt<T>(lf<T>(we[j]) * lf<T>(lt<T>(lf<T>(e[j]) * rs)));
      yr[idx] = *reinterpret_cast<const uint4*>(o);
    }
  }
}

// Rotary embedding in place, HF convention (rotate_half), every operation rounded to the model dtype like the expression
// x * cos + rotate_half(x) * sin evaluated by PyTorch:  x (tokens, heads, hd), cos / sin (seqlen, hd), position = token % seqlen.
// One thread per 8 elements of the lower half and their partners in the upper half: 2 bytes read + 2 written per element.
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(T* __restrict__ x, const T* __restrict__ cs, const T* __restrict__ sn, long long tokens,
                                                   int heads, int hd, int seqlen) {
  const int half_v = hd >> 4;                                  // 16-byte vectors per half head
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = tokens * heads * half_v;
  if (idx >= total) return;
  const int vv = (int)(idx % half_v);
  const long long th = idx / half_v;                            // token * heads + head
  const long long tok = th / heads;
  const int pos = (int)(tok % seqlen);
  T* lo = x + th * hd + vv * 8;
  T* hi = lo + (hd >> 1);
  const uint4 a = *reinterpret_cast<const uint4*>(lo), b = *reinterpret_cast<const uint4*>(hi);
  const uint4 c0 = *reinterpret_cast<const uint4*>(cs + (size_t)pos * hd + vv * 8), c1 = *reinterpret_cast<const uint4*>(cs + (size_t)pos * hd + (hd >> 1) + vv * 8);
  const uint4 s0 = *reinterpret_cast<const uint4*>(sn + (size_t)pos * hd + vv * 8), s1 = *reinterpret_cast<const uint4*>(sn + (size_t)pos * hd + (hd >> 1) + vv * 8);
  const T *ae = reinterpret_cast<const T*>(&a), *be = reinterpret_cast<const T*>(&b);
  const T *c0e = reinterpret_cast<const T*>(&c0), *c1e = reinterpret_cast<const T*>(&c1);
  const T *s0e = reinterpret_cast<const T*>(&s0), *s1e = reinterpret_cast<const T*>(&s1);
  __align__(16) T ol[8], oh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    // lower half: x_lo * cos + (-x_hi) * sin;  upper half: x_hi * cos + x_lo * sin
    const float xl = lf<T>(ae[j]), xh = lf<T>(be[j]);
    const float t1 = lf<T>(lt<T>(xl * lf<T>(c0e[j]))), t2 = lf<T>(lt<T>(-xh * lf<T>(s0e[j])));
    const float t3 = lf<T>(lt<T>(xh * lf<T>(c1e[j]))), t4 = lf<T>(lt<T>(xl * lf<T>(s1e[j])));
    ol[j] = lt<T>(t1 + t2);
    oh[j] = lt<T>(t3 + t4);
  }
  *reinterpret_cast<uint4*>(lo) = *reinterpret_cast<const uint4*>(ol);
  *reinterpret_cast<uint4*>(hi) = *reinterpret_cast<const uint4*>(oh);
}

// out = T(T(silu(g)) * u),  silu evaluated in fp32 and rounded once (F.silu on a 16-bit tensor), the product in the model
// dtype: 4 bytes read + 2 written per element.
template <typename T>
__global__ void __launch_bounds__(256) silu_mul_kernel(const T* g, const T* u, T* out, long long nvec) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nvec) return;
  const uint4 gv = reinterpret_cast<const uint4*>(g)[idx], uv = reinterpret_cast<const uint4*>(u)[idx];
  const T *ge = reinterpret_cast<const T*>(&gv), *ue = reinterpret_cast<const T*>(&uv);
  __align__(16) T o[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float gf = lf<T>(ge[j]);
    const float s = gf / (1.0f + expf(-gf));
    o[j] = lt<T>(lf<T>(lt<T>(s)) * lf<T>(ue[j]));
  }
  reinterpret_cast<uint4*>(out)[idx] = *reinterpret_cast<const uint4*>(o);
}

template <typename T>
int rms_norm_t(const void* x, const void* w, void* y, long long rows, int D, float eps, cudaStream_t st) {
  const int nvec = D >> 3;
  const int vpl = (nvec + 31) / 32;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  GQ_LAUNCHED();
#define GQ_RMS_CASE(V) rms_norm_kernel<T, V><<<grid, 256, 0, st>>>((const T*)x, (const T*)w, (T*)y, rows, D, eps)
  if (vpl <= 4) GQ_RMS_CASE(4);
  else if (vpl <= 8) GQ_RMS_CASE(8);
  else if (vpl <= 16) GQ_RMS_CASE(16);
  else if (vpl <= 32) GQ_RMS_CASE(32);
  else return GQ_ERR_SHAPE;                                    // rows beyond 8192 elements: not a Llama hidden size
#undef GQ_RMS_CASE
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

int llama_rms_norm(const void* x, const void* w, void* y, long long rows, int D, float eps, int fmt, cudaStream_t st) {
  if (rows <= 0 || D <= 0) return GQ_OK;
  if (D & 7) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return GQ_ERR_ALIGN;
  if (fmt == 0) return rms_norm_t<__half>(x, w, y, rows, D, eps, st);
  if (fmt == 1) return rms_norm_t<__nv_bfloat16>(x, w, y, rows, D, eps, st);
  return GQ_ERR_DTYPE;
}

int llama_rope(void* x, const void* cos_t, const void* sin_t, long long tokens, int heads, int hd, int seqlen, int fmt, cudaStream_t st) {
  if (tokens <= 0 || heads <= 0) return GQ_OK;
  if (hd < 16 || (hd & 15) || seqlen <= 0) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(cos_t) | reinterpret_cast<uintptr_t>(sin_t)) & 15) return GQ_ERR_ALIGN;
  const long long total = tokens * heads * (hd >> 4);
  const unsigned grid = (unsigned)((total + 255) / 256);
  GQ_LAUNCHED();
  if (fmt == 0) rope_kernel<__half><<<grid, 256, 0, st>>>((__half*)x, (const __half*)cos_t, (const __half*)sin_t, tokens, heads, hd, seqlen);
  else if (fmt == 1) rope_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)x, (const __nv_bfloat16*)cos_t, (const __nv_bfloat16*)sin_t, tokens, heads, hd, seqlen);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int llama_silu_mul(const void* g, const void* u, void* out, long long n, int fmt, cudaStream_t st) {
  if (n <= 0) return GQ_OK;
  if (n & 7) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(u) | reinterpret_cast<uintptr_t>(out)) & 15) return GQ_ERR_ALIGN;
  const long long nvec = n >> 3;
  const unsigned grid = (unsigned)((nvec + 255) / 256);
  GQ_LAUNCHED();
  if (fmt == 0) silu_mul_kernel<__half><<<grid, 256, 0, st>>>((const __half*)g, (const __half*)u, (__half*)out, nvec);
  else if (fmt == 1) silu_mul_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)g, (const __nv_bfloat16*)u, (__nv_bfloat16*)out, nvec);
  else return GQ_ERR_DTYPE;
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
