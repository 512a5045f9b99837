// K6  W4A16 dequant-fused GEMM on tcgen05:   y[m][n] = sum_k x[m][perm[k]] * (q[n][k] - z[n][g(k)]) * s[n][g(k)]
//
// Replaces reconstruct_nn_linear + nn.Linear (quant.py:262-279, llama.py:172-177), which materialises a dense 16-bit
// W_hat and calls cuBLAS.  One kernel serves every batch size:
//   * the packed int4 weights are the ONLY large stream: they are read exactly once per 64-row batch tile with TMA
//     (128-byte rows, SWIZZLE_128B) into an 8-stage ring of 16 KB boxes = 128 KB in flight per SM (Little: 6.5 TB/s x
//     ~1.5 us / 148 SMs = 66 KB are needed to keep HBM3e busy)
//   * 16 dequant warps turn each landed box into the 16-bit tile W_hat = (q - z) * s with exactly the reference's
//     rounding (integer subtract is exact, ONE 16-bit rounding of the product, quant.py:236) and hand it to the tensor
//     core THROUGH TENSOR MEMORY (tcgen05.st, the MMA's A operand is read from TMEM): a thread owns one weight row =
//     one TMEM lane, so no shuffles and no shared-memory round trip.  (Writing the 4x larger 16-bit tile to shared
//     memory and letting the MMA read it back costs 10 B of smem traffic per packed byte: 230 B/clk/SM at HBM speed,
//     above the 128 B/clk an SM has -- measured at 22% of the HBM roofline before this change.)
//     int4 -> fp16 by the 0x6400 magic number (two codes per LOP3), -> bf16 by 0x4300
//   * the weights are the M = 128 side of tcgen05.mma (kind::f16, fp32 accumulation in TMEM), the batch is the N side
//     (16..64 columns), so batch 1..64 costs the same weight stream and the same dequant work
//   * activations are gathered through the act-order permutation ONCE per call into a scratch (M, K) matrix whose
//     columns are additionally shuffled inside every group of 8 to the order the nibble extraction produces
//     (c0 c4 c1 c5 c2 c6 c3 c7) -- a dot product does not care about a permutation applied to both operands
//   * scale / zero are stored group-major, (G, N): the metadata of the 128 rows of a tile is one coalesced line
//   * work unit = (128 weight rows) x (k chunk); units are dealt round-robin to one persistent CTA per SM; the k chunk
//     is chosen on the host so that the last wave is full; partial sums of split units go through fp32 atomics.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <type_traits>
#include "gptq_b200_internal.h"
#include "ptx.cuh"
#include "w4a16_common.cuh"

namespace gq {
namespace {

constexpr int WM_TILE_N = 128;            // weight rows per unit (UMMA M)
constexpr int WM_KSTAGE = 256;            // k per TMA box of packed weights (128 B per row)
constexpr int WM_KHALF = 128;             // k per dequant / MMA step (one A tile: 2 swizzle atoms of 64)
constexpr int WM_AS = 4;                  // dequantised A tiles in tensor memory (64 columns each)
constexpr int WM_PK_BYTES = WM_TILE_N * 128;
constexpr int WM_A_COLS = WM_KHALF / 2;   // 32-bit TMEM columns of one A tile (two 16-bit values per column)
constexpr int WM_DQ_WARPS = 16;
// 24 warps: 0 weight TMA, 1 MMA, 2 TMEM alloc, 3 activation TMA, 4-19 dequant, 20-23 epilogue
constexpr int WM_THREADS = (4 + WM_DQ_WARPS + 4) * 32;

// Two shapes of the same kernel.  A tcgen05.mma with M = 128, K = 16 holds the tensor core for ~172 clk whatever its N
// (tools/micro/umma_small_n.cu), and every pass over the batch dequantises the weights again, so large batches want the
// widest N: BIG runs 256 batch columns per pass (one fp32 accumulator of 256 TMEM columns + 4 A tiles = all 512
// columns; 2 x 64 KB activation stages), the small shape keeps 64 columns with two accumulators and deeper rings.
template <bool BIG> struct WmCfg {
  static constexpr int PS = BIG ? 4 : 8;                // packed-weight ring stages (16 KB each)
  static constexpr int XS = BIG ? 2 : 4;                // activation tile ring (MPAD * 256 B each)
  static constexpr int MPAD_MAX = BIG ? 256 : 64;       // batch columns per pass (UMMA N)
  static constexpr int NACC = BIG ? 1 : 2;              // fp32 accumulators in tensor memory
  static constexpr int X_BYTES_MAX = MPAD_MAX * WM_KHALF * 2;
  static constexpr int SMEM = PS * WM_PK_BYTES + XS * X_BYTES_MAX + 1024 + 512;
  static constexpr int TMEM_D = 0;                      // accumulators: NACC x MPAD_MAX columns
  static constexpr int TMEM_A = NACC * MPAD_MAX;        // A tiles: WM_AS x WM_A_COLS columns
  static_assert(TMEM_A + WM_AS * WM_A_COLS <= 512, "tensor memory");
  static_assert(SMEM <= 227 * 1024, "shared memory");
};

struct W4Params {
  const void* scale;        // (G, N) 16-bit, group-major: the 128 rows of a tile are one contiguous 256-byte line
  const uint8_t* qzero;     // (G, N)
  void* y;                  // (M, N) 16-bit, written directly when units are not split
  float* yacc;              // (M, N) fp32 zero-initialised accumulator for split units
  int M, N, K, G, group;
  int m_base;               // first batch row of this pass
  int mpad;                 // UMMA N: a multiple of 16, <= 64 (small shape) or <= 256 (BIG)
  int k_chunk;              // k per unit (multiple of 256)
  int n_tiles, k_chunks;
  int split;                // 1: atomics into yacc, 0: direct store
  int group_shift;          // log2(group)
};

template <typename XT, bool BIG>
__global__ void __launch_bounds__(WM_THREADS, 1)
w4a16_mma_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, const W4Params p) {
  using Cfg = WmCfg<BIG>;
  constexpr int WM_PS = Cfg::PS, WM_XS = Cfg::XS, WM_MPAD_MAX = Cfg::MPAD_MAX, WM_X_BYTES_MAX = Cfg::X_BYTES_MAX;
  constexpr int WM_TMEM_D = Cfg::TMEM_D, WM_TMEM_A = Cfg::TMEM_A, NACC = Cfg::NACC;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t pk_base = smem_base;                                  // [PS][128 rows][128 B]  packed codes
  const uint32_t x_base = pk_base + WM_PS * WM_PK_BYTES;               // [XS][2 atoms][mpad rows][128 B] activations
  const uint32_t bar_base = x_base + WM_XS * WM_X_BYTES_MAX;
  auto pk_full = [&](int s) { return bar_base + 8u * s; };
  auto pk_empty = [&](int s) { return bar_base + 8u * (WM_PS + s); };
  auto a_full = [&](int s) { return bar_base + 8u * (2 * WM_PS + s); };
  auto a_empty = [&](int s) { return bar_base + 8u * (2 * WM_PS + WM_AS + s); };
  auto x_full = [&](int s) { return bar_base + 8u * (2 * WM_PS + 2 * WM_AS + s); };
  auto x_empty = [&](int s) { return bar_base + 8u * (2 * WM_PS + 2 * WM_AS + WM_XS + s); };
  auto tfull = [&](int s) { return bar_base + 8u * (2 * WM_PS + 2 * WM_AS + 2 * WM_XS + s); };
  auto tempty = [&](int s) { return bar_base + 8u * (2 * WM_PS + 2 * WM_AS + 2 * WM_XS + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * WM_PS + 2 * WM_AS + 2 * WM_XS + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = p.n_tiles * p.k_chunks;
  const uint32_t x_bytes = (uint32_t)p.mpad * WM_KHALF * 2;
  // number of 128-k steps of unit u (the last chunk may be short)
  auto unit_steps = [&](int u, int& n0, int& k0) {
    const int nt = u / p.k_chunks, kc = u - nt * p.k_chunks;
    n0 = nt * WM_TILE_N;
    k0 = kc * p.k_chunk;
    const int kend = min(p.K, k0 + p.k_chunk);
    return (kend - k0 + WM_KHALF - 1) / WM_KHALF;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_x);
    for (int s = 0; s < WM_PS; ++s) { mbar_init(pk_full(s), 1); mbar_init(pk_empty(s), WM_DQ_WARPS); }
    for (int s = 0; s < WM_AS; ++s) { mbar_init(a_full(s), WM_DQ_WARPS); mbar_init(a_empty(s), 1); }
    for (int s = 0; s < WM_XS; ++s) { mbar_init(x_full(s), 1); mbar_init(x_empty(s), 1); }
    for (int s = 0; s < NACC; ++s) { mbar_init(tfull(s), 1); mbar_init(tempty(s), 4); }
    fence_mbar_init();
  }
  if (warp == 2) { tmem_alloc<1>(tmem_slot, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== packed-weight producer: one 128 x 256-code box per two steps =====================
    if (lane == 0) {
      int ps = 0; uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        for (int st = 0; st < steps; st += 2) {
          mbar_wait_backoff(pk_empty(ps), ph ^ 1u);
          mbar_expect_tx(pk_full(ps), WM_PK_BYTES);
          tma_load_2d(pk_base + ps * WM_PK_BYTES, &map_w, pk_full(ps), (k0 + st * WM_KHALF) / 8, n0);
          if (++ps == WM_PS) { ps = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== activation producer: [mpad][128 k] per step, two 64-k swizzle atoms =====================
    if (lane == 0) {
      int xs = 0; uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        for (int st = 0; st < steps; ++st) {
          mbar_wait_backoff(x_empty(xs), ph ^ 1u);
          mbar_expect_tx(x_full(xs), x_bytes);
          const uint32_t dst = x_base + xs * WM_X_BYTES_MAX;
          tma_load_2d(dst, &map_x, x_full(xs), k0 + st * WM_KHALF, p.m_base);
          tma_load_2d(dst + x_bytes / 2, &map_x, x_full(xs), k0 + st * WM_KHALF + 64, p.m_base);
          if (++xs == WM_XS) { xs = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(std::is_same<XT, __nv_bfloat16>::value ? 1u : 0u, 0u, 0u,
                                        WM_TILE_N, (uint32_t)p.mpad);
      int as = 0, xs = 0; uint32_t aph = 0, xph = 0;
      int it = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        const int acc = it % NACC;
        mbar_wait(tempty(acc), ((uint32_t)(it / NACC) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + WM_TMEM_D + acc * WM_MPAD_MAX;
        for (int st = 0; st < steps; ++st) {
          mbar_wait(a_full(as), aph);
          mbar_wait(x_full(xs), xph);
          tc_fence_after();
          const uint32_t ta = tmem_base + WM_TMEM_A + as * WM_A_COLS;
          const uint32_t sx = x_base + xs * WM_X_BYTES_MAX;
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t bd = kmajor_sw128_desc(sx + h * (x_bytes / 2) + kk * 32);
              mma_f16_ts(d_tmem, ta + (h * 4 + kk) * 8, bd, idesc, (st | h | kk) ? 1u : 0u);   // 16 k = 8 columns
            }
          tc_commit_cg1(a_empty(as));
          tc_commit_cg1(x_empty(xs));
          if (++as == WM_AS) { as = 0; aph ^= 1u; }
          if (++xs == WM_XS) { xs = 0; xph ^= 1u; }
        }
        tc_commit_cg1(tfull(acc));
      }
    }
  } else if (warp >= 4 && warp < 4 + WM_DQ_WARPS) {
    // ===================== dequant: packed codes -> W_hat tile in TENSOR MEMORY (A operand) =====================
    // A warp may touch TMEM lanes 32 (warp % 4) .. +31 only: thread = weight row (lane of TMEM), the four warps of a
    // lane quarter split the 128 k of a step into 32-k parts (one 16-byte chunk of packed codes, 16 TMEM columns).
    const int quarter = warp & 3, kpart = (warp - 4) >> 2;
    const int row = quarter * 32 + lane;
    const int r7 = row & 7;
    const int group_shift = p.group_shift;
    int ps = 0, as = 0; uint32_t pph = 0, aph = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      int n0, k0;
      const int steps = unit_steps(u, n0, k0);
      const int n = min(n0 + row, p.N - 1);
      // Out-of-range rows / groups are clamped instead of predicated: such rows are never stored and such columns meet
      // zero-filled activations, so whatever finite value comes back is harmless.  The RAW 16-bit / 8-bit values are
      // fetched one step ahead and only converted when used, so the load latency hides behind a whole step.
      const uint16_t* sp = reinterpret_cast<const uint16_t*>(p.scale) + n;
      const uint8_t* zp = p.qzero + n;
      auto group_of = [&](int st) { return min((k0 + st * WM_KHALF + kpart * 32) >> group_shift, p.G - 1); };
      uint16_t sc_n = sp[(size_t)group_of(0) * p.N];
      uint8_t zz_n = zp[(size_t)group_of(0) * p.N];
      for (int st = 0; st < steps; ++st) {
        const int half_in_box = st & 1;
        const uint16_t sc = sc_n;
        const uint8_t zz = zz_n;
        if (st + 1 < steps) {
          const size_t go = (size_t)group_of(st + 1) * p.N;
          sc_n = sp[go]; zz_n = zp[go];
        }
        if (half_in_box == 0) mbar_wait(pk_full(ps), pph);
        const uint8_t* pk = smem_gen + (pk_base - smem_base) + ps * WM_PK_BYTES + row * 128;
        const int chunk = half_in_box * 4 + kpart;            // 16-byte chunk (32 codes) of this row's 128-byte line
        const uint4 wv = *reinterpret_cast<const uint4*>(pk + ((chunk ^ r7) << 4));
        Dq<XT> dq;
        dq.set(sc, (uint32_t)zz);
        uint32_t v[16];
        {
          const uint4 o0 = dq.cvt(wv.x), o1 = dq.cvt(wv.y), o2 = dq.cvt(wv.z), o3 = dq.cvt(wv.w);
          v[0] = o0.x; v[1] = o0.y; v[2] = o0.z; v[3] = o0.w; v[4] = o1.x; v[5] = o1.y; v[6] = o1.z; v[7] = o1.w;
          v[8] = o2.x; v[9] = o2.y; v[10] = o2.z; v[11] = o2.w; v[12] = o3.x; v[13] = o3.y; v[14] = o3.z; v[15] = o3.w;
        }
        mbar_wait(a_empty(as), aph ^ 1u);
        tc_fence_after();
        tmem_st_32x16(tmem_base + (uint32_t(quarter * 32) << 16) + WM_TMEM_A + as * WM_A_COLS + kpart * 16, v);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(a_full(as));
          if (half_in_box == 1 || st + 1 == steps) mbar_arrive(pk_empty(ps));
        }
        if (half_in_box == 1 || st + 1 == steps) { if (++ps == WM_PS) { ps = 0; pph ^= 1u; } }
        if (++as == WM_AS) { as = 0; aph ^= 1u; }
      }
    }
  } else if (warp >= 4 + WM_DQ_WARPS) {
    // ===================== epilogue: TMEM -> y (or fp32 atomics for split units) =====================
    const int we = warp & 3;                          // TMEM lane quarter
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
      int n0, k0;
      unit_steps(u, n0, k0);
      const int acc = it % NACC;
      mbar_wait_backoff(tfull(acc), (uint32_t)(it / NACC) & 1u);
      tc_fence_after();
      const int n = n0 + 32 * we + lane;
      const uint32_t t_row = tmem_base + (uint32_t(32 * we) << 16) + WM_TMEM_D + acc * WM_MPAD_MAX;
      for (int c = 0; c < p.mpad; c += 16) {
        uint32_t v[16];
        tmem_ld_32x16(t_row + c, v);
        tmem_ld_wait();
        if (c + 16 >= p.mpad) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty(acc));
        }
        if (n < p.N) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int m = p.m_base + c + i;
            if (m < p.M) {
              const float val = __uint_as_float(v[i]);
              if (p.split) atomicAdd(p.yacc + (size_t)m * p.N + n, val);
              else reinterpret_cast<XT*>(p.y)[(size_t)m * p.N + n] = cvt_out<XT>(val);
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<1>(tmem_base, 512);
}

template <typename XT, bool BIG>
int run_w4a16_mma_cfg(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
                      void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  using Cfg = WmCfg<BIG>;
  constexpr int WM_MPAD_MAX = Cfg::MPAD_MAX, WM_SMEM = Cfg::SMEM;
  static PerDeviceFlag attr_set;   // function attributes are per device
  static int n_sms = 0;
  if (!attr_set) {
    if (cudaFuncSetAttribute(w4a16_mma_kernel<XT, BIG>, cudaFuncAttributeMaxDynamicSharedMemorySize, WM_SMEM) != cudaSuccess)
      return GQ_ERR_CUDA;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
    attr_set = true;
  }
  // workspace: [shuffled activations (M x K) 16-bit][fp32 accumulator (M x N)]
  XT* xs = reinterpret_cast<XT*>(workspace);
  const size_t xs_bytes = (((size_t)M * K * 2) + 255) & ~(size_t)255;
  float* yacc = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + xs_bytes);
  dim3 pg((K + 255) / 256, M);
  GQ_LAUNCHED();
  w4a16_xshuffle_kernel<XT><<<pg, 256, 0, st>>>(reinterpret_cast<const XT*>(x), perm, xs, M, K);

  // k chunk: the split that minimises (waves x steps per unit + 1 step of fill/drain per unit)
  const int n_tiles = (N + WM_TILE_N - 1) / WM_TILE_N;
  const int boxes = (K + WM_KSTAGE - 1) / WM_KSTAGE;
  int best_split = 1;
  double best_cost = 1e30;
  for (int sp = 1; sp <= boxes && sp <= 64; ++sp) {
    const int per = (boxes + sp - 1) / sp;
    const int chunks = (boxes + per - 1) / per;
    const long units = (long)n_tiles * chunks;
    const long waves = (units + n_sms - 1) / n_sms;
    const double cost = (double)waves * (per + 1.0) + (chunks > 1 ? 0.5 : 0.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best_split = chunks; }
  }
  const int per = (boxes + best_split - 1) / best_split;
  const int k_chunks = (boxes + per - 1) / per;

  CUtensorMap map_w;
  int rc = make_tmap_2d_raw(&map_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT32, (uint64_t)(K / 8), (uint64_t)N, (uint64_t)(K / 8) * 4,
                            32, WM_TILE_N, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != GQ_OK) return rc;
  const int split = k_chunks > 1;
  if (split && cudaMemsetAsync(yacc, 0, (size_t)M * N * sizeof(float), st) != cudaSuccess) return GQ_ERR_CUDA;
  for (int m_base = 0; m_base < M; m_base += WM_MPAD_MAX) {
    const int mrem = M - m_base;
    const int mpad = mrem >= WM_MPAD_MAX ? WM_MPAD_MAX : ((mrem + 15) / 16) * 16;
    CUtensorMap map_x;
    rc = make_tmap_2d_raw(&map_x, xs, fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                          (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, 64, mpad, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != GQ_OK) return rc;
    W4Params p;
    p.scale = scale; p.qzero = qzero; p.y = y; p.yacc = yacc;
    p.M = M; p.N = N; p.K = K; p.G = K / group; p.group = group;
    p.m_base = m_base; p.mpad = mpad; p.k_chunk = per * WM_KSTAGE; p.n_tiles = n_tiles; p.k_chunks = k_chunks; p.split = split;
    p.group_shift = 0;
    while ((1 << p.group_shift) < group) ++p.group_shift;
    const long units = (long)n_tiles * k_chunks;
    const int grid = (int)(units < n_sms ? units : n_sms);
    GQ_LAUNCHED();
    w4a16_mma_kernel<XT, BIG><<<grid, WM_THREADS, WM_SMEM, st>>>(map_w, map_x, p);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  }
  if (split) {
    const size_t n = (size_t)M * N;
    GQ_LAUNCHED();
    w4a16_finalize_kernel<XT><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(yacc, reinterpret_cast<XT*>(y), n);
  }
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

template <typename XT>
int run_w4a16_mma(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
                  void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M > WmCfg<false>::MPAD_MAX) return run_w4a16_mma_cfg<XT, true>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
  return run_w4a16_mma_cfg<XT, false>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
}

}  // namespace

size_t w4a16_workspace_bytes(int M, int N, int K) {
  // xs (M, K) 16-bit | max(yacc (M, N) fp32 of the tcgen05 split-K path, chunk sums (M, K/128) fp32 of the small / mid batch paths)
  const size_t tail = (size_t)M * sizeof(float) * (size_t)(N > K / 128 + 8 ? N : K / 128 + 8);
  return ((((size_t)M * K * 2) + 255) & ~(size_t)255) + tail;
}

int w4a16_mma(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
               void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M <= 0 || N <= 0) return GQ_OK;
  if (group < 32 || (group & (group - 1)) != 0 || K % group != 0) return GQ_ERR_SHAPE;   // group: power of two >= 32
  if ((reinterpret_cast<uintptr_t>(qw) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 255)) return GQ_ERR_ALIGN;
  if (!workspace) return GQ_ERR_ARG;
  if (fmt == 0) return run_w4a16_mma<__half>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
  if (fmt == 1) return run_w4a16_mma<__nv_bfloat16>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
  return GQ_ERR_DTYPE;
}

// Batch <= 8 goes to the warp-level mma kernel (w4a16_smallm.cu), batch <= 16 to its windowed form (w4a16_midm.cu), batch
// 1-2 on shapes those refuse to the CUDA-core GEMV (w4a16_gemv.cu), everything else to the tcgen05 kernel: measured on B200, a
// tcgen05.mma with M = 128, K = 16 occupies the tensor core for 171.6 cycles for EVERY N from 16 to 256
// (tools/micro/umma_small_n.cu), so with the weights on the M side the tensor path streams at most
// 128*16/2 B / 171.6 clk * 148 SMs * 1.85 GHz = 1.6 TB/s of packed int4 however small the batch is.
int w4a16_gemm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm, void* y,
               void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (g_w4a16_path == 2 && M <= 128 && w4a16_wb_supports(M, N, K, group))      // experimental weights-on-N kernel, opt-in
    return w4a16_wb(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
  if (g_w4a16_path == 4 && w4a16_midm_supports(M, N, K, group))                 // A/B: the windowed kernel wherever it runs
    return w4a16_midm(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
  if ((g_w4a16_path == 0 || g_w4a16_path == 3) && w4a16_smallm_supports(M, N, K, group))
    return w4a16_smallm(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
  // batch 9..16, and batch 3..8 on rows too long for the small-batch kernel's resident activation: its windowed form where
  // it measured faster than the tcgen05 path (profiles/r02_v38_w4a16_midm_sweep.txt): a CTA needs several row tiles, and
  // with two blocks of batch columns a K that is long for the number of rows (4096 x 11008) still favours the split over K
  if (g_w4a16_path == 0 && M > 2 && (M <= 8 ? N >= 2048 : (N >= 8192 || (N >= 4096 && K <= 8192))) &&
      w4a16_midm_supports(M, N, K, group))
    return w4a16_midm(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
  if (M <= 2 && K <= 16384) return w4a16_gemv(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
  return w4a16_mma(x, qw, scale_t, qzero_t, perm, y, workspace, M, N, K, group, fmt, st);
}

}  // namespace gq
