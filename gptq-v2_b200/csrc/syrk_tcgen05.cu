// K1  Hessian accumulation  H += X^T X   (upper-triangular SYRK, fp32 accumulate)
//
// Replaces the reference's Triton kernel (gptq_triton/accumulate_hessian.py:55-149)
// and its CUTLASS sm80 GEMM (package/gptq/accumulate_hessian_kernel.cu:10-84).
//
// Design (sm_100a only):
//   * X is (T, D) row-major 16-bit.  Both MMA operands are tiles of the SAME matrix
//     and both are "MN-major" (the K index = token is the strided one):
//         H[m][n] += sum_t X[t][m] * X[t][n]
//   * Work unit = 256x256 super-tile (a, nj) of H with nj >= a (upper triangle only).
//     CG==2: a CTA pair (cluster of 2) computes it with tcgen05.mma.cta_group::2
//            (UMMA 256x256x16): CTA r stages A rows [128r,128r+128) and HALF of B,
//            so each SM pulls 32 KB from L2 per 64-token k-block (128 flop/B).
//     CG==1: each CTA computes one 128x256 half on its own (UMMA 128x256x16).
//   * TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) brings [64 tokens][64 columns]
//     sub-blocks (8 KB) into a multi-stage smem ring; one thread issues the MMAs;
//     fp32 accumulators live in TMEM, double buffered (2 x 256 columns) so the
//     epilogue of tile i overlaps the main loop of tile i+1.
//   * Epilogue: tcgen05.ld -> registers -> read-modify-write of H in global memory.
//       mode 0: upper triangle only (col >= row)        -- HessianHook fast path
//       mode 1: upper RMW + mirrored store to the lower -- Triton semantics
//       mode 2: RMW of both triangles                   -- CUTLASS semantics (H may be asymmetric)
//   * Persistent: one CTA (pair) per SM (pair), static round-robin over super-tiles in an L2-friendly order
//     (decode_tile: 8 x 8 blocks of super-tiles).
#include "ptx.cuh"
#include "gptq_b200_internal.h"

namespace gq {

template <int CG> struct SyrkCfg {
  static constexpr int BK = 64;                      // tokens per k-block
  static constexpr int SUB_BYTES = 64 * 128;         // one [64 tok][64 col] 16-bit sub-block
  static constexpr int A_SUBS = 2;                   // 128 rows of H per CTA
  static constexpr int B_SUBS = (CG == 2) ? 2 : 4;   // 128 (half) or 256 columns of H per CTA
  static constexpr int STAGE_BYTES = (A_SUBS + B_SUBS) * SUB_BYTES;
  static constexpr int STAGES = (CG == 2) ? 6 : 4;
  static constexpr int UMMA_M = 128 * CG;
  static constexpr int UMMA_N = 256;
  static constexpr int TMEM_COLS = 512;              // 2 accumulator buffers x 256 fp32 columns
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

struct SyrkParams {
  float* H;
  int D;
  int T;
  int mode;
  int fmt;        // 0 fp16, 1 bf16
  int NT;         // ceil(D / 256)
  int n_tiles;    // NT (NT + 1) / 2
  int kchunk;     // 64-token k-blocks accumulated in tensor memory before the partial sum is added to H (0 = all of T)
  int groups;     // D % 64 == 0: operands through the 3D column-group map (one TMA instruction per operand and k-block)
};

// Tile order.  Super-tiles that run at the same time should share operand panels: the 74 CTA pairs walk the upper
// triangle in blocks of RASTER x RASTER super-tiles (row-major over blocks, row-major inside a block), so one wave
// touches ~2 x RASTER..3 x RASTER distinct 256-column panels of X instead of ~NT (a plain row-major walk of the
// triangle: 12.8 GB of DRAM reads per launch at D = 11008, 10x the algorithmic bytes).
constexpr int RASTER = 8;
constexpr int RASTER_MIN_NT = 24;                      // below this a wave covers most panels anyway: plain row-major walk
__device__ __forceinline__ void decode_tile(int s, int NT, int& a, int& nj) {
  if (NT < RASTER_MIN_NT) {
    a = 0;
    int rowlen = NT;
    while (s >= rowlen) { s -= rowlen; ++a; --rowlen; }
    nj = a + s;
    return;
  }
  const int nblk = (NT + RASTER - 1) / RASTER;
  for (int bi = 0; bi < nblk; ++bi) {
    const int h = min(RASTER, NT - bi * RASTER);
    // diagonal block: upper triangle of h x h
    const int cd = h * (h + 1) / 2;
    if (s < cd) {
      int row = 0, rowlen = h;
      while (s >= rowlen) { s -= rowlen; ++row; --rowlen; }
      a = bi * RASTER + row;
      nj = a + s;
      return;
    }
    s -= cd;
    for (int bj = bi + 1; bj < nblk; ++bj) {
      const int w = min(RASTER, NT - bj * RASTER);
      const int c = h * w;
      if (s < c) {
        const int row = s / w;
        a = bi * RASTER + row;
        nj = bj * RASTER + (s - row * w);
        return;
      }
      s -= c;
    }
  }
  a = 0; nj = 0;                                       // not reached for s < NT (NT + 1) / 2
}

template <int CG>
__global__ void __launch_bounds__(256, 1)
syrk_upper_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap3, const SyrkParams p) {
  using C = SyrkCfg<CG>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES;
  // barrier layout (8 B each): full[STAGES] | empty[STAGES] | tfull[2] | tempty[2] | tmem_ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + C::STAGES * C::STAGE_BYTES +
                                                                          8 * (2 * C::STAGES + 4));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = (CG == 2) ? cluster_ctarank() : (blockIdx.x & 1u);
  const int pair = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;
  const int num_kb = (p.T + C::BK - 1) / C::BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap); tma_prefetch_desc(&tmap3);
    for (int s = 0; s < C::STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 4 * CG); }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<CG>(tmem_slot, C::TMEM_COLS);
    tmem_relinquish<CG>();
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs of a pair) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int s = pair; s < p.n_tiles; s += n_pairs) {
        int a, nj; decode_tile(s, p.NT, a, nj);
        const int m_base = a * 256 + 128 * (int)rank;
        const int n_base = nj * 256 + ((CG == 2) ? 128 * (int)rank : 0);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          const int t0 = kb * C::BK;
          if constexpr (CG == 2) {
            // all bytes of both CTAs are accounted on the leader's barrier
            const uint32_t lead_full = full_bar(stage) & 0xFEFFFFFFu;
            if (rank == 0) mbar_expect_tx(full_bar(stage), 2u * C::STAGE_BYTES);
            if (p.groups) {
              // X as (64 columns, tokens, column groups): both 64-column sub-tiles of an operand in one instruction
              tma_load_3d_cg2(sa, &tmap3, lead_full, 0, t0, m_base >> 6);
              tma_load_3d_cg2(sa + C::A_SUBS * C::SUB_BYTES, &tmap3, lead_full, 0, t0, n_base >> 6);
            } else {
#pragma unroll
            for (int i = 0; i < C::A_SUBS; ++i)
              tma_load_2d_cg2(sa + i * C::SUB_BYTES, &tmap, lead_full, m_base + 64 * i, t0);
#pragma unroll
            for (int j = 0; j < C::B_SUBS; ++j)
              tma_load_2d_cg2(sa + (C::A_SUBS + j) * C::SUB_BYTES, &tmap, lead_full, n_base + 64 * j, t0);
            }
          } else {
            mbar_expect_tx(full_bar(stage), C::STAGE_BYTES);
#pragma unroll
            for (int i = 0; i < C::A_SUBS; ++i)
              tma_load_2d(sa + i * C::SUB_BYTES, &tmap, full_bar(stage), m_base + 64 * i, t0);
#pragma unroll
            for (int j = 0; j < C::B_SUBS; ++j)
              tma_load_2d(sa + (C::A_SUBS + j) * C::SUB_BYTES, &tmap, full_bar(stage), n_base + 64 * j, t0);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only when CG==2) =====================
    if ((CG == 1 || rank == 0) && lane == 0) {
      const uint32_t idesc = umma_idesc((uint32_t)p.fmt, 1u, 1u, C::UMMA_M, C::UMMA_N);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      // The tensor core adds every MMA into the fp32 accumulator with truncation: over the 2048 MMAs of a 32768-token
      // call the sums come out ~5e-6 low (relative), which the inverse factor amplifies by the condition number.  The
      // token range therefore leaves tensor memory in chunks; the epilogue adds each partial sum to H in fp32
      // round-to-nearest while the next chunk accumulates in the other TMEM buffer.
      const int kstep = p.kchunk > 0 ? p.kchunk : num_kb;
      for (int s = pair; s < p.n_tiles; s += n_pairs)
      for (int kc = 0; kc < num_kb; kc += kstep, ++it) {
        const int kce = min(kc + kstep, num_kb);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = kc; kb < kce; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sb = sa + C::A_SUBS * C::SUB_BYTES;
#pragma unroll
          for (int k4 = 0; k4 < C::BK / 16; ++k4) {
            // MN-major SW128: 16 tokens = 2 swizzle atoms of 8 rows x 128 B -> +2048 B per MMA
            const uint64_t ad = umma_smem_desc_sw128(sa + k4 * 2048, C::SUB_BYTES, 1024);
            const uint64_t bd = umma_smem_desc_sw128(sb + k4 * 2048, C::SUB_BYTES, 1024);
            mma_f16<CG>(d_tmem, ad, bd, idesc, ((kb - kc) | k4) ? 1u : 0u);
          }
          if constexpr (CG == 2) tc_commit_cg2_mcast(empty_bar(stage), 0x3); else tc_commit_cg1(empty_bar(stage));
          if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
        if constexpr (CG == 2) tc_commit_cg2_mcast(tfull_bar(acc), 0x3); else tc_commit_cg1(tfull_bar(acc));
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> H (RMW) =====================
    const int we = warp - 4;                       // TMEM lane quarter this warp may access
    const int D = p.D;
    int it = 0;
    const int kstep = p.kchunk > 0 ? p.kchunk : num_kb;
    for (int s = pair; s < p.n_tiles; s += n_pairs)
    for (int kc = 0; kc < num_kb; kc += kstep, ++it) {
      int a, nj; decode_tile(s, p.NT, a, nj);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1u;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int row = a * 256 + 128 * (int)rank + 32 * we + lane;
      const bool diag = (nj == a);
      const bool full_direct = !diag || (p.mode == 2);      // every element of the tile is wanted
      const bool transposed_rmw = (p.mode == 2) && !diag;
      const bool mirror = (p.mode == 1);
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
          if (lane == 0) {
            if constexpr (CG == 2) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
            else mbar_arrive(tempty_bar(acc));
          }
        }
        const int col0 = nj * 256 + c * 32;
        if (row < D && col0 < D) {
          float* hrow = p.H + (size_t)row * D + col0;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const int col = col0 + 4 * q;
            if (col < D && (full_direct || col + 3 >= row)) {
              float4 h = *reinterpret_cast<float4*>(hrow + 4 * q);
              float n0 = h.x + __uint_as_float(v[4 * q + 0]);
              float n1 = h.y + __uint_as_float(v[4 * q + 1]);
              float n2 = h.z + __uint_as_float(v[4 * q + 2]);
              float n3 = h.w + __uint_as_float(v[4 * q + 3]);
              if (full_direct || col + 0 >= row) h.x = n0;
              if (full_direct || col + 1 >= row) h.y = n1;
              if (full_direct || col + 2 >= row) h.z = n2;
              if (full_direct || col + 3 >= row) h.w = n3;
              *reinterpret_cast<float4*>(hrow + 4 * q) = h;
              if (mirror) {
                v[4 * q + 0] = __float_as_uint(n0); v[4 * q + 1] = __float_as_uint(n1);
                v[4 * q + 2] = __float_as_uint(n2); v[4 * q + 3] = __float_as_uint(n3);
              }
            }
          }
          if (mirror) {
            // lanes = consecutive rows -> each store instruction is one coalesced 128 B line
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int col = col0 + i;
              if (col < D && col > row) p.H[(size_t)col * D + row] = __uint_as_float(v[i]);
            }
          } else if (transposed_rmw) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int col = col0 + i;
              if (col < D) p.H[(size_t)col * D + row] += __uint_as_float(v[i]);
            }
          }
        }
      }
    }
  }

  // ---- teardown
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, C::TMEM_COLS);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(ptr);
  }
  return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_kind, uint64_t inner, uint64_t outer,
                 uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, int swizzle_atom32) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return GQ_ERR_DRIVER;
  CUtensorMapDataType dt = elem_kind == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                         : elem_kind == 1 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                          : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? GQ_OK : GQ_ERR_TENSORMAP;
}

// fp32 matrix (rows x cols, cols % 32 == 0) seen as (32 columns, rows, cols / 32 column groups): one box of (32, box_rows,
// box_groups) lands as box_groups consecutive [box_rows][32] sub-tiles (SWIZZLE_128B_ATOM_32B, the MN-major tf32 operand
// layout) -- one TMA instruction where a 2D map needs one per 32-column group.
int make_tmap_f32_colgroups(CUtensorMap* out, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                            uint32_t box_groups) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return GQ_ERR_DRIVER;
  if (cols % 32 != 0 || ld % 4 != 0) return GQ_ERR_SHAPE;
  cuuint64_t dims[3] = {32, rows, cols / 32};
  cuuint64_t strides[2] = {ld * 4, 128};
  cuuint32_t box[3] = {32, box_rows, box_groups};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? GQ_OK : GQ_ERR_TENSORMAP;
}

int make_tmap_2d_raw(CUtensorMap* out, const void* base, int cu_dtype, uint64_t inner, uint64_t outer,
                     uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, int cu_swizzle) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return GQ_ERR_DRIVER;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, static_cast<CUtensorMapDataType>(cu_dtype), 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, static_cast<CUtensorMapSwizzle>(cu_swizzle),
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? GQ_OK : GQ_ERR_TENSORMAP;
}

template <int CG>
static int launch_syrk(const CUtensorMap& tmap, const CUtensorMap& tmap3, const SyrkParams& p, int n_sms, cudaStream_t st) {
  using C = SyrkCfg<CG>;
  static PerDeviceFlag attr_set;   // function attributes are per device
  if (!attr_set) {
    if (cudaFuncSetAttribute(syrk_upper_kernel<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES) !=
        cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  int pairs = n_sms / 2;
  if (pairs > p.n_tiles) pairs = p.n_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (CG == 2) ? 2 : 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  GQ_LAUNCHED();
  cudaError_t e = cudaLaunchKernelEx(&cfg, syrk_upper_kernel<CG>, tmap, tmap3, p);
  return e == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int syrk_upper(float* H, const void* X, int D, int T, int fmt, int mode, int variant, cudaStream_t st) {
  if (D <= 0 || T <= 0) return GQ_OK;
  if (D % 8 != 0) return GQ_ERR_SHAPE;            // TMA needs 16-byte row pitch
  if (fmt != 0 && fmt != 1) return GQ_ERR_DTYPE;
  if ((reinterpret_cast<uintptr_t>(X) & 15) || (reinterpret_cast<uintptr_t>(H) & 15)) return GQ_ERR_ALIGN;
  CUtensorMap tmap;
  int rc = make_tmap_2d(&tmap, X, fmt, (uint64_t)D, (uint64_t)T, (uint64_t)D * 2, 64, 64, 0);
  if (rc != GQ_OK) return rc;
  SyrkParams p;
  p.H = H; p.D = D; p.T = T; p.mode = mode; p.fmt = fmt;
  p.NT = (D + 255) / 256;
  p.n_tiles = p.NT * (p.NT + 1) / 2;
  p.kchunk = g_syrk_kchunk;
  p.groups = (D % 64 == 0) ? 1 : 0;
  CUtensorMap tmap3 = tmap;
  if (p.groups) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) return GQ_ERR_DRIVER;
    cuuint64_t dims[3] = {64, (cuuint64_t)T, (cuuint64_t)(D / 64)};
    cuuint64_t strides[2] = {(cuuint64_t)D * 2, 128};
    cuuint32_t box[3] = {64, 64, 2};
    cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&tmap3, fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(X), dims,
            strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return GQ_ERR_TENSORMAP;
  }
  int dev = 0, n_sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
  if (variant == 1) return launch_syrk<1>(tmap, tmap3, p, n_sms, st);
  return launch_syrk<2>(tmap, tmap3, p, n_sms, st);
}

}  // namespace gq
