// Pieces shared by the two tcgen05 W4A16 kernels (w4a16_mma.cu: weights on the M side through tensor memory;
// w4a16_wb.cu: weights on the N side through shared memory): activation pre-shuffle, fp32 -> 16-bit finalisation,
// the K-major SWIZZLE_128B descriptor and the packed word -> 8 dequantised values conversion.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "gptq_b200_internal.h"
#include "ptx.cuh"

namespace gq {
namespace {

template <typename XT> __device__ __forceinline__ XT cvt_out(float v);
template <> __device__ __forceinline__ __half cvt_out<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// xs[m][8 j + 2 i + e] = x[m][perm[8 j + i + 4 e]]   (i = 0..3, e = 0..1)
template <typename XT>
__global__ void w4a16_xshuffle_kernel(const XT* __restrict__ x, const int32_t* __restrict__ perm, XT* __restrict__ xs,
                                      int M, int K) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y;
  if (idx >= K) return;
  const int j = idx >> 3, s = idx & 7, i = s >> 1, e = s & 1;
  const int k = 8 * j + i + 4 * e;
  const int pk = perm ? perm[k] : k;
  xs[(size_t)m * K + idx] = x[(size_t)m * K + pk];
}

template <typename XT>
__global__ void w4a16_finalize_kernel(const float* __restrict__ yacc, XT* __restrict__ y, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = cvt_out<XT>(yacc[i]);
}

// K-major SWIZZLE_128B shared-memory descriptor: 8-row groups of 128 B rows are 1024 B apart (SBO); LBO unused.
__device__ __forceinline__ uint64_t kmajor_sw128_desc(uint32_t saddr) {
  return umma_smem_desc(saddr, 16, 1024, 2u);
}

// (a & mask) | magic as ONE LOP3 (the C expression compiles to two because both constants are immediates)
template <uint32_t MASK, uint32_t MAGIC>
__device__ __forceinline__ uint32_t and_or(uint32_t a) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "n"(MASK), "n"(MAGIC));
  return r;
}

// one packed word (8 codes) -> 8 dequantised 16-bit values in the order c0 c4 c1 c5 c2 c6 c3 c7
template <typename XT> struct Dq;
template <> struct Dq<__half> {
  __half2 zl, zh, s2;
  __device__ __forceinline__ void set(uint16_t s_bits, uint32_t z) {
    const uint32_t zlo = 0x64006400u + z * 0x00010001u;                 // 1024 + z (exact)
    zl = *reinterpret_cast<const __half2*>(&zlo);
    const uint32_t zhi = 0xD400D400u + z * 0x00100010u;                 // -(64 + z): -64 = 0xD400, ulp 1/16 -> z << 4
    zh = *reinterpret_cast<const __half2*>(&zhi);                       // for the x16 nibbles: (1024 + 16 q) / 16 - 64 - z
    const uint32_t sb = (uint32_t)s_bits * 0x00010001u;
    s2 = *reinterpret_cast<const __half2*>(&sb);
  }
  __device__ __forceinline__ uint4 cvt(uint32_t w) const {
    const __half2 sixteenth = __float2half2_rn(0.0625f);
    const uint32_t w8 = w >> 8;
    uint32_t b0 = and_or<0x000F000Fu, 0x64006400u>(w);      // codes 0,4
    uint32_t b1 = and_or<0x00F000F0u, 0x64006400u>(w);      // codes 1,5 (x16)
    uint32_t b2 = and_or<0x000F000Fu, 0x64006400u>(w8);     // codes 2,6
    uint32_t b3 = and_or<0x00F000F0u, 0x64006400u>(w8);     // codes 3,7 (x16)
    __half2 d0 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&b0), zl), s2);
    __half2 d1 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&b1), sixteenth, zh), s2);
    __half2 d2 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&b2), zl), s2);
    __half2 d3 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&b3), sixteenth, zh), s2);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&d0); o.y = *reinterpret_cast<uint32_t*>(&d1);
    o.z = *reinterpret_cast<uint32_t*>(&d2); o.w = *reinterpret_cast<uint32_t*>(&d3);
    return o;
  }
};
template <> struct Dq<__nv_bfloat16> {
  __nv_bfloat162 zl, s2;
  __device__ __forceinline__ void set(uint16_t s_bits, uint32_t z) {
    const uint32_t zlo = 0x43004300u + z * 0x00010001u;                 // 128 + z (exact in bf16 for z < 128)
    zl = *reinterpret_cast<const __nv_bfloat162*>(&zlo);
    const uint32_t sb = (uint32_t)s_bits * 0x00010001u;
    s2 = *reinterpret_cast<const __nv_bfloat162*>(&sb);
  }
  __device__ __forceinline__ uint4 cvt(uint32_t w) const {
    uint32_t b0 = and_or<0x000F000Fu, 0x43004300u>(w);
    uint32_t b1 = and_or<0x000F000Fu, 0x43004300u>(w >> 4);
    uint32_t b2 = and_or<0x000F000Fu, 0x43004300u>(w >> 8);
    uint32_t b3 = and_or<0x000F000Fu, 0x43004300u>(w >> 12);
    __nv_bfloat162 d0 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162*>(&b0), zl), s2);
    __nv_bfloat162 d1 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162*>(&b1), zl), s2);
    __nv_bfloat162 d2 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162*>(&b2), zl), s2);
    __nv_bfloat162 d3 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162*>(&b3), zl), s2);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&d0); o.y = *reinterpret_cast<uint32_t*>(&d1);
    o.z = *reinterpret_cast<uint32_t*>(&d2); o.w = *reinterpret_cast<uint32_t*>(&d3);
    return o;
  }
};
}  // namespace
}  // namespace gq
