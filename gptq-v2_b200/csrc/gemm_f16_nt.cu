// Dense 16-bit "NT" GEMM on tcgen05 CTA pairs:   y (M x N) = x (M x K) * w (N x K)^T,   fp32 accumulation in TMEM.
//
// The large-batch half of K6.  The reference evaluates a quantised linear as a dense 16-bit reconstruction + cuBLAS
// (quant.py:262-279 + nn.Linear, llama.py:172-177); for batches that are compute-bound this build dequantises the
// packed weights ONCE into a transient 16-bit matrix (gq_w4a16_dequant, the reference's rounding) and multiplies
// with this kernel, so that no library GEMM is left on the product path (round 1 called torch.matmul here).
//
// Same skeleton as K1 (syrk_tcgen05.cu): persistent CTA pairs (cluster of 2), tcgen05.mma.cta_group::2 256 x 256 x 16,
// a 6-stage TMA ring, double-buffered fp32 accumulators in tensor memory so that the epilogue of tile i overlaps the main
// loop of tile i + 1.  Both operands are K-major here (x rows and weight rows are contiguous along k): one
// [64 k][128 rows] SWIZZLE_128B box per operand, CTA and k-block (CTA r stages its 128 rows of x and its half of the 256
// weight rows: 32 KB per 64-deep k-block and SM).  Tiles are walked m-fastest inside a 256-row weight panel, so a panel
// is read from HBM once and then served by L2 to the CTA pairs working on the same panel.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "gptq_b200_internal.h"
#include "ptx.cuh"

namespace gq {
namespace {

constexpr int GN_BK = 64;                          // k per stage (one 128-byte swizzle row of 16-bit values)
constexpr int GN_TILE_BYTES = 128 * 128;           // [128 rows][64 k] 16-bit
constexpr int GN_STAGE_BYTES = 2 * GN_TILE_BYTES;  // this CTA's x rows + its half of the weight rows
constexpr int GN_STAGES = 6;
constexpr int GN_SMEM_BYTES = GN_STAGES * GN_STAGE_BYTES + 1024 + 256;

struct GemmNtParams {
  void* y;
  int M, N, K;
  int fmt;                 // 0 fp16, 1 bf16
  int tiles_m, tiles_n;
};

template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// K-major SWIZZLE_128B: 8-row groups of 128-byte rows are 1024 B apart (SBO); LBO unused
__device__ __forceinline__ uint64_t gn_desc(uint32_t saddr) { return umma_smem_desc(saddr, 16, 1024, 2u); }

template <typename T>
__global__ void __launch_bounds__(256, 1)
gemm_f16_nt_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                   const GemmNtParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + GN_STAGES * GN_STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (GN_STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * GN_STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * GN_STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * GN_STAGES + 4);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_gen + GN_STAGES * GN_STAGE_BYTES + 8 * (2 * GN_STAGES + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int num_kb = (p.K + GN_BK - 1) / GN_BK;
  const int n_tiles = p.tiles_m * p.tiles_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_w);
    for (int s = 0; s < GN_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 8); }
    fence_mbar_init();
  }
  if (warp == 2) { tmem_alloc<2>(tmem_slot, 512); tmem_relinquish<2>(); }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs of the pair) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = pair; t < n_tiles; t += n_pairs) {
        const int tn = t / p.tiles_m, tm = t - tn * p.tiles_m;        // m fastest inside a weight panel
        const int m_base = tm * 256 + 128 * (int)rank, n_base = tn * 256 + 128 * (int)rank;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * GN_STAGE_BYTES;
          // all bytes of both CTAs are accounted on the leader's barrier
          const uint32_t lead_full = full_bar(stage) & 0xFEFFFFFFu;
          if (rank == 0) mbar_expect_tx(full_bar(stage), 2u * GN_STAGE_BYTES);
          tma_load_2d_cg2(sa, &map_x, lead_full, kb * GN_BK, m_base);
          tma_load_2d_cg2(sa + GN_TILE_BYTES, &map_w, lead_full, kb * GN_BK, n_base);
          if (++stage == GN_STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA) =====================
    if (rank == 0 && lane == 0) {
      const uint32_t idesc = umma_idesc((uint32_t)p.fmt, 0u, 0u, 256, 256);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * GN_STAGE_BYTES;
          const uint32_t sb = sa + GN_TILE_BYTES;
#pragma unroll
          for (int kk = 0; kk < GN_BK / 16; ++kk)                      // 16 k = 32 bytes inside the swizzle row
            mma_f16<2>(d_tmem, gn_desc(sa + kk * 32), gn_desc(sb + kk * 32), idesc, (kb | kk) ? 1u : 0u);
          tc_commit_cg2_mcast(empty_bar(stage), 0x3);
          if (++stage == GN_STAGES) { stage = 0; phase ^= 1u; }
        }
        tc_commit_cg2_mcast(tfull_bar(acc), 0x3);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> 16-bit -> y =====================
    const int we = warp - 4;                          // TMEM lane quarter this warp may access
    T* y = reinterpret_cast<T*>(p.y);
    int it = 0;
    for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
      const int tn = t / p.tiles_m, tm = t - tn * p.tiles_m;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = tm * 256 + 128 * (int)rank + 32 * we + lane;
      const uint32_t t_row = tmem_base + (uint32_t(32 * we) << 16) + acc * 256;
#pragma unroll 1
      for (int c = 0; c < 8; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(t_row + c * 32, v);
        tmem_ld_wait();
        if (c == 7) {
          // accumulator buffer fully read by this warp -> hand it back to the MMA issuer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
        }
        const int col0 = tn * 256 + c * 32;
        if (row < p.M && col0 < p.N) {
          T* dst = y + (size_t)row * p.N + col0;
          if (col0 + 32 <= p.N && (p.N & 7) == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack2<T>(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
              o.y = pack2<T>(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack2<T>(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack2<T>(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              *reinterpret_cast<uint4*>(dst + 8 * q) = o;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              if (col0 + i + 1 < p.N) {
                *reinterpret_cast<uint32_t*>(dst + i) = pack2<T>(__uint_as_float(v[i]), __uint_as_float(v[i + 1]));
              } else if (col0 + i < p.N) {
                const uint32_t pk = pack2<T>(__uint_as_float(v[i]), 0.f);
                dst[i] = *reinterpret_cast<const T*>(&pk);
              }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc<2>(tmem_base, 512);
}

template <typename T>
int launch_gemm_nt(const CUtensorMap& mx, const CUtensorMap& mw, const GemmNtParams& p, int n_sms, cudaStream_t st) {
  static PerDeviceFlag attr_set;   // function attributes are per device
  if (!attr_set) {
    if (cudaFuncSetAttribute(gemm_f16_nt_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, GN_SMEM_BYTES) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  int pairs = n_sms / 2;
  const int tiles = p.tiles_m * p.tiles_n;
  if (pairs > tiles) pairs = tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = GN_SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  GQ_LAUNCHED();
  return cudaLaunchKernelEx(&cfg, gemm_f16_nt_kernel<T>, mx, mw, p) == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

// y (M, N) = x (M, K) w (N, K)^T; x, w, y of the same 16-bit type, row-major contiguous, 16-byte aligned; K % 8 == 0,
// N % 2 == 0 (full-speed stores need N % 8 == 0).
int gemm_f16_nt(const void* x, const void* w, void* y, int M, int N, int K, int fmt, cudaStream_t st) {
  if (M <= 0 || N <= 0) return GQ_OK;
  if (fmt != 0 && fmt != 1) return GQ_ERR_DTYPE;
  if (K <= 0 || (K & 7) || (N & 1)) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y)) & 15) return GQ_ERR_ALIGN;
  CUtensorMap mx, mw;
  int rc;
  if ((rc = make_tmap_2d(&mx, x, fmt, (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, GN_BK, 128, 0)) != GQ_OK) return rc;
  if ((rc = make_tmap_2d(&mw, w, fmt, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, GN_BK, 128, 0)) != GQ_OK) return rc;
  GemmNtParams p;
  p.y = y; p.M = M; p.N = N; p.K = K; p.fmt = fmt;
  p.tiles_m = (M + 255) / 256;
  p.tiles_n = (N + 255) / 256;
  int dev = 0, n_sms = 0;
  cudaGetDevice(&dev);
  if (cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sms <= 0) n_sms = 148;
  if (fmt == 0) return launch_gemm_nt<__half>(mx, mw, p, n_sms, st);
  return launch_gemm_nt<__nv_bfloat16>(mx, mw, p, n_sms, st);
}

}  // namespace gq
