// K6  W4A16 dequant-fused GEMM, batch 1-2 path on the CUDA cores (w4a16_mma.cu serves batch >= 3: a tcgen05.mma
// costs 171.6 cycles per 128 rows x 16 k whatever the batch width, which caps the tensor path at ~1.5 TB/s of int4).
//
//   y[m][n] = sum_k x[m][perm[k]] * (q[n][k] - z[n][g(k)]) * s[n][g(k)]
//
// replaces reconstruct_nn_linear + nn.Linear (quant.py:262-279, llama.py:172-177), which
// materialises a dense fp16 W_hat and calls cuBLAS.  The packed weights (0.5 B/element) are
// streamed exactly once with 16-byte non-allocating loads; everything else is arranged so
// that the weight stream is the only significant traffic:
//   * w4a16_xprep_kernel gathers x through the act-order permutation ONCE into a scratch
//     buffer laid out as the nibble-extraction order needs it: for every 32-column chunk u
//     and word wq, four half2 pairs (c0,c4)(c1,c5)(c2,c6)(c3,c7)  (64 B per chunk per row of x)
//   * in the main kernel a CTA owns a block of weight rows and its warps split K into
//     2048-column slices: a lane keeps the activations of its two chunk positions in registers
//     and re-uses them for every row of the block; 8 weight loads per lane are in flight
//   * int4 -> fp16 by the 0x6400 magic-number trick, two codes per LOP3; the nibble at bits
//     4..7 is converted with one HFMA2 (x/16 - 64 - z) instead of shift + HSUB2
//   * fp16 HFMA2 over 32 codes, then fp32 accumulation with the group scale.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "../../include/gptq_b200.h"
#include "gptq_b200_internal.h"

namespace gq {
namespace {

template <typename XT> __device__ __forceinline__ __half x_to_half(XT v);
template <> __device__ __forceinline__ __half x_to_half<__half>(__half v) { return v; }
template <> __device__ __forceinline__ __half x_to_half<__nv_bfloat16>(__nv_bfloat16 v) {
  return __float2half_rn(fminf(fmaxf(__bfloat162float(v), -65504.f), 65504.f));
}
template <typename XT> __device__ __forceinline__ float s_to_f32(XT v);
template <> __device__ __forceinline__ float s_to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float s_to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename XT> __device__ __forceinline__ XT y_from_f32(float v);
template <> __device__ __forceinline__ __half y_from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 y_from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// xprep[(m * K/32 + u) * 4 + wq] (uint4) = 4 half2 for word wq of chunk u
template <typename XT>
__global__ void w4a16_xprep_kernel(const XT* __restrict__ x, const int32_t* __restrict__ perm,
                                   uint4* __restrict__ xprep, int M, int K) {
  const int wi = blockIdx.x * blockDim.x + threadIdx.x;     // word index along K
  const int m = blockIdx.y;
  if (wi >= K / 8) return;
  __half h[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = wi * 8 + i;
    const int pk = perm ? perm[k] : k;
    h[i] = x_to_half<XT>(x[(size_t)m * K + pk]);
  }
  uint4 o;
  __half2 p0 = __halves2half2(h[0], h[4]), p1 = __halves2half2(h[1], h[5]);
  __half2 p2 = __halves2half2(h[2], h[6]), p3 = __halves2half2(h[3], h[7]);
  o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
  o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
  xprep[((size_t)m * (K / 32) + (wi >> 2)) * 4 + (wi & 3)] = o;
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

constexpr int GEMV_WARPS = 4;
constexpr int SLICE_CHUNKS = 64;      // a K-slice = 64 chunks of 32 columns = 2048 columns (2 chunks per lane)

// One CTA owns RPW consecutive weight rows; its 4 warps take the K-slices round-robin, so a
// lane keeps the activations of its two chunk positions in registers (32 regs per row of x) and
// re-uses them for all RPW rows.  Weight loads are issued 8 x 16 B per lane at a time (4 rows x
// 2 positions) to keep enough bytes in flight at <= ~128 registers per thread.
template <typename XT, int MT, int RPW>
__global__ void __launch_bounds__(GEMV_WARPS * 32) w4a16_gemv_kernel(
    const uint4* __restrict__ xprep, const uint32_t* __restrict__ qw, const XT* __restrict__ scale,
    const uint8_t* __restrict__ qzero, XT* __restrict__ y, int M, int N, int K, int group, int m0) {
  __shared__ float red[GEMV_WARPS][RPW * MT];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_base = blockIdx.x * RPW;
  const int G = K / group;
  const int n_chunks = K / 32;
  const int wpr = K / 8;
  const int n_slices = (n_chunks + SLICE_CHUNKS - 1) / SLICE_CHUNKS;
  float acc[RPW][MT];
#pragma unroll
  for (int i = 0; i < RPW; ++i)
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[i][m] = 0.f;
  const __half2 sixteenth = __float2half2_rn(0.0625f);

  for (int slice = warp; slice < n_slices; slice += GEMV_WARPS) {
    int u[2];
    bool live[2];
    int g[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      u[p] = slice * SLICE_CHUNKS + p * 32 + lane;
      live[p] = u[p] < n_chunks;
      if (!live[p]) u[p] = 0;
      g[p] = (u[p] * 32) / group;
    }
    uint4 xr[MT][2][4];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
      for (int p = 0; p < 2; ++p)
#pragma unroll
        for (int wq = 0; wq < 4; ++wq)
          xr[m][p][wq] = (live[p] && m0 + m < M) ? xprep[((size_t)(m0 + m) * n_chunks + u[p]) * 4 + wq]
                                                  : make_uint4(0, 0, 0, 0);
#pragma unroll 1
    for (int rg = 0; rg < RPW / 4; ++rg) {
      uint4 wv[4][2];
      float sc[4][2];
      uint32_t zz[4][2];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int n = min(n_base + rg * 4 + i, N - 1);     // clamped rows are computed but never stored
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          wv[i][p] = ldg_stream(reinterpret_cast<const uint4*>(qw + (size_t)n * wpr) + u[p]);
          sc[i][p] = live[p] ? s_to_f32<XT>(scale[(size_t)g[p] * N + n]) : 0.f;   // group-major (G, N)
          zz[i][p] = qzero[(size_t)g[p] * N + n];
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          // low nibble of a byte: (1024 + q) - (1024 + z);  high nibble: (1024 + 16 q)/16 - (64 + z)
          const uint32_t zlo = 0x64006400u + zz[i][p] * 0x00010001u;
          const __half2 zl = *reinterpret_cast<const __half2*>(&zlo);
          const __half2 zh = __float2half2_rn(-64.f - (float)zz[i][p]);
          const uint32_t ww[4] = {wv[i][p].x, wv[i][p].y, wv[i][p].z, wv[i][p].w};
          __half2 a2[MT][2];
#pragma unroll
          for (int m = 0; m < MT; ++m) { a2[m][0] = __float2half2_rn(0.f); a2[m][1] = a2[m][0]; }
#pragma unroll
          for (int wq = 0; wq < 4; ++wq) {
            const uint32_t w0 = ww[wq], w8 = ww[wq] >> 8;
            uint32_t b0 = (w0 & 0x000F000Fu) | 0x64006400u;     // codes 0,4
            uint32_t b1 = (w0 & 0x00F000F0u) | 0x64006400u;     // codes 1,5 (x16)
            uint32_t b2 = (w8 & 0x000F000Fu) | 0x64006400u;     // codes 2,6
            uint32_t b3 = (w8 & 0x00F000F0u) | 0x64006400u;     // codes 3,7 (x16)
            __half2 qd[4];
            qd[0] = __hsub2(*reinterpret_cast<__half2*>(&b0), zl);
            qd[1] = __hfma2(*reinterpret_cast<__half2*>(&b1), sixteenth, zh);
            qd[2] = __hsub2(*reinterpret_cast<__half2*>(&b2), zl);
            qd[3] = __hfma2(*reinterpret_cast<__half2*>(&b3), sixteenth, zh);
#pragma unroll
            for (int m = 0; m < MT; ++m) {
              const __half2* xp = reinterpret_cast<const __half2*>(&xr[m][p][wq]);
#pragma unroll
              for (int t = 0; t < 4; ++t) a2[m][wq & 1] = __hfma2(qd[t], xp[t], a2[m][wq & 1]);
            }
          }
#pragma unroll
          for (int m = 0; m < MT; ++m) {
            const float2 f0 = __half22float2(a2[m][0]);
            const float2 f1 = __half22float2(a2[m][1]);
            // acc is indexed with a compile-time row inside the group and a runtime group: keep it static
            const float part = sc[i][p] * ((f0.x + f0.y) + (f1.x + f1.y));
#pragma unroll
            for (int r = 0; r < RPW / 4; ++r)
              if (r == rg) acc[r * 4 + i][m] += part;
          }
        }
      }
    }
  }
  // ---- reduce over lanes, then over the 4 warps (K-slices)
#pragma unroll
  for (int i = 0; i < RPW; ++i)
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      float v = acc[i][m];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][i * MT + m] = v;
    }
  __syncthreads();
  for (int e = threadIdx.x; e < RPW * MT; e += GEMV_WARPS * 32) {
    const int i = e / MT, m = e % MT;
    const float v = (red[0][e] + red[1][e]) + (red[2][e] + red[3][e]);
    if (n_base + i < N && m0 + m < M) y[(size_t)(m0 + m) * N + n_base + i] = y_from_f32<XT>(v);
  }
}

template <typename XT, int MT, int RPW>
void launch_gemv(const uint4* xprep, const uint32_t* qw, const XT* scale, const uint8_t* qzero, XT* y, int M, int N,
                 int K, int group, int m0, cudaStream_t st) {
  GQ_LAUNCHED();
  w4a16_gemv_kernel<XT, MT, RPW><<<(N + RPW - 1) / RPW, GEMV_WARPS * 32, 0, st>>>(xprep, qw, scale, qzero, y, M, N, K,
                                                                                   group, m0);
}

template <typename XT>
int run_w4a16(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
              void* workspace, int M, int N, int K, int group, cudaStream_t st) {
  const XT* xx = reinterpret_cast<const XT*>(x);
  const XT* ss = reinterpret_cast<const XT*>(scale);
  XT* yy = reinterpret_cast<XT*>(y);
  uint4* xprep = reinterpret_cast<uint4*>(workspace);
  dim3 pg((K / 8 + 127) / 128, M);
  GQ_LAUNCHED();
  w4a16_xprep_kernel<XT><<<pg, 128, 0, st>>>(xx, perm, xprep, M, K);
  // rows per CTA: more rows amortise the activation registers, fewer rows give more CTAs
  const bool big = (size_t)N * K >= (size_t)48 * 1024 * 1024;
  int m0 = 0;
  while (m0 < M) {
    const int rem = M - m0;
    if (rem >= 2) {
      if (big) launch_gemv<XT, 2, 16>(xprep, qw, ss, qzero, yy, M, N, K, group, m0, st);
      else launch_gemv<XT, 2, 8>(xprep, qw, ss, qzero, yy, M, N, K, group, m0, st);
      m0 += 2;
    } else {
      if (big) launch_gemv<XT, 1, 32>(xprep, qw, ss, qzero, yy, M, N, K, group, m0, st);
      else launch_gemv<XT, 1, 16>(xprep, qw, ss, qzero, yy, M, N, K, group, m0, st);
      m0 += 1;
    }
  }
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

size_t w4a16_gemv_workspace_bytes(int M, int K) { return (size_t)M * K * 2; }

int w4a16_gemv(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
               void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M <= 0 || N <= 0) return GQ_OK;
  if (group <= 0 || group % 32 != 0 || K % group != 0 || K % 32 != 0) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qw) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 15)) return GQ_ERR_ALIGN;
  if (!workspace) return GQ_ERR_ARG;
  if (fmt == 0) return run_w4a16<__half>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, st);
  if (fmt == 1) return run_w4a16<__nv_bfloat16>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, st);
  return GQ_ERR_DTYPE;
}

// ---- allocation-free batch 1-2 variant (gq_w4a16_gemv_static, see include/gptq_b200.h) --------------------------------
// The only scratch of this path is the shuffled copy of the activations (M * K 16-bit values).  Here it lives in a
// file-scope device array, so a caller that may not allocate (stream capture, a serving loop, a harness that forbids
// cudaMalloc) passes inputs and outputs only.  Every GPU that loads the library holds these 64 KB until exit.
// w4a16_xprep_kernel rewrites all M * K values the GEMV reads on every call, so nothing has to be cleared between calls;
// all calls share the array, so calls that overlap on one GPU (two streams, two threads) race on it.
constexpr int GEMV_STATIC_MAX_M = 2;
constexpr int GEMV_STATIC_MAX_K = 16384;
static __device__ __align__(256) unsigned char g_gemv_static_scratch[GEMV_STATIC_MAX_M * GEMV_STATIC_MAX_K * 2];

namespace {
// The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
// device current, and cached.
std::atomic<void*> g_gemv_static_ptr[GQ_MAX_DEVICES];

int gemv_static_scratch(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= GQ_MAX_DEVICES) return GQ_ERR_CUDA;
  void* ptr = g_gemv_static_ptr[dev].load(std::memory_order_acquire);
  if (!ptr) {
    if (cudaGetSymbolAddress(&ptr, g_gemv_static_scratch) != cudaSuccess || !ptr) return GQ_ERR_CUDA;
    g_gemv_static_ptr[dev].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return GQ_OK;
}
}  // namespace

int w4a16_gemv_static(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm,
                      void* y, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M < 0 || N < 0 || K <= 0) return GQ_ERR_ARG;
  if (M > GEMV_STATIC_MAX_M || K > GEMV_STATIC_MAX_K) return GQ_ERR_LIMIT;
  if (M == 0 || N == 0) return GQ_OK;
  if (!x || !qw || !scale || !qzero || !y) return GQ_ERR_ARG;
  void* scratch = nullptr;
  const int rc = gemv_static_scratch(&scratch);
  if (rc != GQ_OK) return rc;
  return w4a16_gemv(x, qw, scale, qzero, perm, y, scratch, M, N, K, group, fmt, st);
}

}  // namespace gq

extern "C" int gq_w4a16_gemv_static(const void* x, const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t,
                                    const int32_t* perm, void* y, int M, int N, int K, int group, int dtype,
                                    void* stream) {
  return gq::w4a16_gemv_static(x, qweight, scale_t, qzero_t, perm, y, M, N, K, group, dtype,
                               reinterpret_cast<cudaStream_t>(stream));
}
