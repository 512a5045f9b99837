// K6 small / mid batch: pieces shared by w4a16_smallm.cu (batch <= 8, whole activation in shared memory) and
// w4a16_midm.cu (batch <= 16, activation streamed through shared memory in windows): the int4 -> 16-bit fragment unpack,
// the warp-level MMA wrappers, the streaming weight load and the activation gather into fragment order.
#pragma once
#include "gptq_b200_internal.h"
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace gq {
namespace {

constexpr int SM_WARPS = 16;                 // one CTA of 512 threads per SM
constexpr int SM_THREADS = SM_WARPS * 32;
constexpr int SM_CHUNK = 128;                // k per unit

template <uint32_t MASK, uint32_t MAGIC>
__device__ __forceinline__ uint32_t and_or(uint32_t a) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "n"(MASK), "n"(MAGIC));
  return r;
}

template <typename T> struct Fmt;
// unpack(w, o): one packed word -> four 2 x 16-bit registers  o0 = (n0, n4), o1 = (n1, n5), o2 = (n2, n6), o3 = (n3, n7),
// each nibble OR-ed into the mantissa of a constant whose ulp lines up with it.  The ALU pipe (LOP3, PRMT, right
// shifts: one warp instruction per 2 clk per scheduler) is what this kernel runs out of first, so the byte permute
// v = (byte 1, 0, byte 3, 0) serves two registers and left shifts are left to the FMA pipe (IMAD.SHL).
template <> struct Fmt<__half> {
  static constexpr uint32_t MASK = 0x00f000f0u, MAGIC = 0x54005400u;   // 64 + q, ulp 1/16: nibble on mantissa bits 4..7
  static constexpr float OFF = 64.f;
  __device__ static __forceinline__ void unpack(uint32_t w, uint32_t (&o)[4]) {
    const uint32_t v = __byte_perm(w, 0u, 0x4341u);
    o[0] = and_or<MASK, MAGIC>(w << 4);
    o[1] = and_or<MASK, MAGIC>(w);
    o[2] = and_or<MASK, MAGIC>(v << 4);
    o[3] = and_or<MASK, MAGIC>(v);
  }
  __device__ static __forceinline__ float to_f(__half v) { return __half2float(v); }
  __device__ static __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  __device__ static __forceinline__ void mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
};
template <> struct Fmt<__nv_bfloat16> {
  static constexpr uint32_t MASK = 0x000f000fu, MAGIC = 0x43004300u;   // 128 + q, ulp 1: nibble on mantissa bits 0..3
  static constexpr float OFF = 128.f;
  __device__ static __forceinline__ void unpack(uint32_t w, uint32_t (&o)[4]) {
    const uint32_t v = __byte_perm(w, 0u, 0x4341u);
    o[0] = and_or<MASK, MAGIC>(w);
    o[1] = and_or<MASK, MAGIC>(w >> 4);
    o[2] = and_or<MASK, MAGIC>(v);
    o[3] = and_or<MASK, MAGIC>(v >> 4);
  }
  __device__ static __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ static __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  __device__ static __forceinline__ void mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint4 ldg_stream(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// One 8-element slot of the fragment order: gathers x[n][perm[...]] for source word `sl` (16 slots per 128-k chunk),
// returns the 16 bytes to store and the slot's sum.
//   slot (c, i, t) of column n holds  x[n][perm[c*128 + t*32 + i*8 + j + 4*e]]  at  2*j + e      (j = 0..3, e = 0..1)
// i.e. lane t's i-th 16 bytes are the B registers of the two MMAs fed by component i of its 16-byte weight load.
template <typename T>
__device__ __forceinline__ void gather_slot(const T* __restrict__ xrow, const int32_t* __restrict__ perm, int k0, uint4& out,
                                            float& sum) {
  int pk[8];
  if (perm) {
    const int4 a = *reinterpret_cast<const int4*>(perm + k0), b = *reinterpret_cast<const int4*>(perm + k0 + 4);
    pk[0] = a.x; pk[1] = a.y; pk[2] = a.z; pk[3] = a.w; pk[4] = b.x; pk[5] = b.y; pk[6] = b.z; pk[7] = b.w;
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) pk[j] = k0 + j;
  }
  T v[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[2 * j] = xrow[pk[j]];
    v[2 * j + 1] = xrow[pk[j + 4]];
  }
  sum = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) sum += Fmt<T>::to_f(v[j]);
  out = *reinterpret_cast<const uint4*>(v);
}

// Large M*K: the gather is done once here instead of once per CTA.  xs_pre[n][c*128 + i*32 + t*8 ..], s_pre[n][c].
template <typename T>
__global__ void w4a16_smallm_xprep_kernel(const T* __restrict__ x, const int32_t* __restrict__ perm, T* __restrict__ xs_pre,
                                          float* __restrict__ s_pre, int K) {
  const int n = blockIdx.y;
  const int sl = blockIdx.x * blockDim.x + threadIdx.x;          // blockDim.x % 32 == 0, K/8 % 16 == 0
  const bool valid = sl < K / 8;
  const int cc = sl >> 4, tt = (sl >> 2) & 3, ii = sl & 3;
  uint4 out = make_uint4(0, 0, 0, 0);
  float sum = 0.f;
  if (valid) {
    gather_slot<T>(x + (size_t)n * K, perm, cc * SM_CHUNK + tt * 32 + ii * 8, out, sum);
    *reinterpret_cast<uint4*>(xs_pre + (size_t)n * K + cc * SM_CHUNK + ii * 32 + tt * 8) = out;
  }
  sum += __shfl_xor_sync(0xffffffffu, sum, 8);
  sum += __shfl_xor_sync(0xffffffffu, sum, 4);
  sum += __shfl_xor_sync(0xffffffffu, sum, 2);
  sum += __shfl_xor_sync(0xffffffffu, sum, 1);
  if (valid && (sl & 15) == 0) s_pre[(size_t)n * (K / SM_CHUNK) + cc] = sum;
}

inline int sm_count() {
  static int n_sms = 0;
  if (n_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sms <= 0) n_sms = 148;
  }
  return n_sms;
}

}  // namespace
}  // namespace gq
