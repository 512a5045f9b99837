// K6, batch <= 128, EXPERIMENTAL (opt-in: gq_set_w4a16_path(2)):  y = x . W_hat^T  with the WEIGHTS ON THE N SIDE of tcgen05.mma.
// Status: parity-green, but 88-187 us at 28672 x 8192 for batch 16..128 against 85-113 us of the weights-on-M kernel: the
// tensor pipe is 36 % active (the 16 dequant warps + fence.proxy.async per 64-k step do not keep up) and the split-k epilogue
// issues one scalar atomic per element with lanes along the batch.  Kept for the next round; not dispatched by default.
//
// A tcgen05.mma with M = 128, K = 16 holds the tensor core for ~172 clk whatever its N (tools/micro/umma_small_n.cu).
// With the weights as the A operand (w4a16_mma.cu) one instruction consumes 128 weight rows x 16 k, which caps the
// weight stream at 1.6 TB/s of packed int4 however small the batch is.  Here the batch (padded to 128 rows by TMA
// zero fill) is A and 256 weight rows are B, so one instruction consumes twice the weights: the cap moves to
// 3.3 TB/s, at the price of writing the dequantised tile to shared memory (B has to be a shared-memory operand).
//
//   * work unit = (256 weight rows) x (k chunk), persistent CTA per SM, split units reduced through fp32 atomics
//   * packed codes: one TMA box [256 rows][64 B] (SWIZZLE_64B) per 128 k, ring of 4 (refills follow consumption closely)
//   * 16 dequant warps: thread = (weight row, 32-k half of a 64-k step): one LDS.128 of packed codes -> four 16-byte
//     chunks of the K-major SWIZZLE_128B B tile [256 rows][64 k] (conflict-free: the 8 rows of a quarter-warp land on
//     8 different chunks), fence.proxy.async, mbarrier; ring of 3 tiles
//   * activations: pre-shuffled copy (w4a16_xshuffle_kernel), TMA box [128 rows][64 k] per step, rows >= M zero-filled
//   * D: 128 lanes (batch) x 256 columns (weight rows) fp32, two accumulators = all 512 TMEM columns, so the
//     epilogue of a unit overlaps the main loop of the next
#include <type_traits>
#include "w4a16_common.cuh"

namespace gq {
namespace {

constexpr int WB_TILE_N = 256;            // weight rows per unit (UMMA N)
constexpr int WB_KBOX = 128;              // k per packed TMA box
constexpr int WB_KSTEP = 64;              // k per dequant / MMA step (one swizzle atom)
constexpr int WB_PS = 4;                  // packed ring stages (16 KB each)
constexpr int WB_BS = 3;                  // dequantised B tiles (32 KB each)
constexpr int WB_XS = 3;                  // activation tiles (16 KB each)
constexpr int WB_PK_BYTES = WB_TILE_N * 64;
constexpr int WB_B_BYTES = WB_TILE_N * WB_KSTEP * 2;
constexpr int WB_X_BYTES = 128 * WB_KSTEP * 2;
constexpr int WB_SMEM = WB_PS * WB_PK_BYTES + WB_BS * WB_B_BYTES + WB_XS * WB_X_BYTES + 1024 + 512;
constexpr int WB_DQ_WARPS = 16;
// 24 warps: 0 weight TMA, 1 MMA, 2 TMEM alloc, 3 activation TMA, 4-19 dequant, 20-23 epilogue
constexpr int WB_THREADS = (4 + WB_DQ_WARPS + 4) * 32;
static_assert(WB_SMEM <= 227 * 1024, "shared memory");

struct WbParams {
  const void* scale;        // (G, N) 16-bit, group-major
  const uint8_t* qzero;     // (G, N)
  void* y;                  // (M, N) 16-bit, written directly when units are not split
  float* yacc;              // (M, N) fp32 zero-initialised accumulator for split units
  int M, N, K, G;
  int m_base;               // first batch row of this pass (passes of 128 rows)
  int k_chunk;              // k per unit (multiple of 128)
  int n_tiles, k_chunks;
  int split;                // 1: atomics into yacc, 0: direct store
  int group_shift;          // log2(group)
};

template <typename XT>
__global__ void __launch_bounds__(WB_THREADS, 1)
w4a16_wb_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, const WbParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t pk_base = smem_base;                                  // [PS][2 boxes][128 rows][128 B] packed codes
  const uint32_t b_base = pk_base + WB_PS * WB_PK_BYTES;               // [BS][256 rows][128 B]  W_hat, K-major SW128
  const uint32_t x_base = b_base + WB_BS * WB_B_BYTES;                 // [XS][128 rows][128 B]  activations
  const uint32_t bar_base = x_base + WB_XS * WB_X_BYTES;
  auto pk_full = [&](int s) { return bar_base + 8u * s; };
  auto pk_empty = [&](int s) { return bar_base + 8u * (WB_PS + s); };
  auto b_full = [&](int s) { return bar_base + 8u * (2 * WB_PS + s); };
  auto b_empty = [&](int s) { return bar_base + 8u * (2 * WB_PS + WB_BS + s); };
  auto x_full = [&](int s) { return bar_base + 8u * (2 * WB_PS + 2 * WB_BS + s); };
  auto x_empty = [&](int s) { return bar_base + 8u * (2 * WB_PS + 2 * WB_BS + WB_XS + s); };
  auto tfull = [&](int s) { return bar_base + 8u * (2 * WB_PS + 2 * WB_BS + 2 * WB_XS + s); };
  auto tempty = [&](int s) { return bar_base + 8u * (2 * WB_PS + 2 * WB_BS + 2 * WB_XS + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * WB_PS + 2 * WB_BS + 2 * WB_XS + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = p.n_tiles * p.k_chunks;
  // number of 64-k steps of unit u (the last chunk may be short; K % 64 == 0)
  auto unit_steps = [&](int u, int& n0, int& k0) {
    const int nt = u / p.k_chunks, kc = u - nt * p.k_chunks;
    n0 = nt * WB_TILE_N;
    k0 = kc * p.k_chunk;
    const int kend = min(p.K, k0 + p.k_chunk);
    return (kend - k0) / WB_KSTEP;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_x);
    for (int s = 0; s < WB_PS; ++s) { mbar_init(pk_full(s), 1); mbar_init(pk_empty(s), WB_DQ_WARPS); }
    for (int s = 0; s < WB_BS; ++s) { mbar_init(b_full(s), WB_DQ_WARPS); mbar_init(b_empty(s), 1); }
    for (int s = 0; s < WB_XS; ++s) { mbar_init(x_full(s), 1); mbar_init(x_empty(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull(s), 1); mbar_init(tempty(s), 4); }
    fence_mbar_init();
  }
  if (warp == 2) { tmem_alloc<1>(tmem_slot, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== packed-weight producer: one [256 rows][128 codes] box per two steps =====================
    if (lane == 0) {
      int ps = 0; uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        for (int st = 0; st < steps; st += WB_KBOX / WB_KSTEP) {
          mbar_wait_backoff(pk_empty(ps), ph ^ 1u);
          mbar_expect_tx(pk_full(ps), WB_PK_BYTES);
          tma_load_2d(pk_base + ps * WB_PK_BYTES, &map_w, pk_full(ps), (k0 + st * WB_KSTEP) / 8, n0);
          if (++ps == WB_PS) { ps = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== activation producer: [128 rows][64 k] per step (rows >= M arrive as zeros) ==============
    if (lane == 0) {
      int xs = 0; uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        for (int st = 0; st < steps; ++st) {
          mbar_wait_backoff(x_empty(xs), ph ^ 1u);
          mbar_expect_tx(x_full(xs), WB_X_BYTES);
          tma_load_2d(x_base + xs * WB_X_BYTES, &map_x, x_full(xs), k0 + st * WB_KSTEP, p.m_base);
          if (++xs == WB_XS) { xs = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: D[batch 128][weight rows 256] += X[128][16] . W_hat[256][16]^T ================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(std::is_same<XT, __nv_bfloat16>::value ? 1u : 0u, 0u, 0u, 128u, (uint32_t)WB_TILE_N);
      int bs = 0, xs = 0; uint32_t bph = 0, xph = 0;
      int it = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
        int n0, k0;
        const int steps = unit_steps(u, n0, k0);
        const int acc = it & 1;
        mbar_wait(tempty(acc), ((it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * WB_TILE_N;
        for (int st = 0; st < steps; ++st) {
          mbar_wait(b_full(bs), bph);
          mbar_wait(x_full(xs), xph);
          tc_fence_after();
          const uint32_t sb = b_base + bs * WB_B_BYTES, sx = x_base + xs * WB_X_BYTES;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_f16<1>(d_tmem, kmajor_sw128_desc(sx + kk * 32), kmajor_sw128_desc(sb + kk * 32), idesc, (st | kk) ? 1u : 0u);
          tc_commit_cg1(b_empty(bs));
          tc_commit_cg1(x_empty(xs));
          if (++bs == WB_BS) { bs = 0; bph ^= 1u; }
          if (++xs == WB_XS) { xs = 0; xph ^= 1u; }
        }
        tc_commit_cg1(tfull(acc));
      }
    }
  } else if (warp >= 4 && warp < 4 + WB_DQ_WARPS) {
    // ===================== dequant: packed codes -> W_hat tile in shared memory (B operand) =====================
    const int dt = threadIdx.x - 128;                 // 0..511
    const int row = dt & 255, khalf = dt >> 8;        // weight row of the tile, 32-k half of the step
    const int r7 = row & 7;
    const int r64 = (row >> 1) & 3;                   // SWIZZLE_64B: 16-byte chunk index ^= address bits [7:8]
    const uint32_t pk_row = (uint32_t)row * 64u;
    const int group_shift = p.group_shift;
    int ps = 0, bs = 0; uint32_t pph = 0, bph = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      int n0, k0;
      const int steps = unit_steps(u, n0, k0);
      const int n = min(n0 + row, p.N - 1);           // rows past N are never stored: clamp instead of predicating
      const uint16_t* sp = reinterpret_cast<const uint16_t*>(p.scale) + n;
      const uint8_t* zp = p.qzero + n;
      auto group_of = [&](int st) { return min((k0 + st * WB_KSTEP + khalf * 32) >> group_shift, p.G - 1); };
      // metadata runs three steps ahead of its use (raw 16-bit / 8-bit values, converted when used)
      auto meta_at = [&](int st, uint16_t& s, uint8_t& z) {
        const size_t go = (size_t)group_of(min(st, steps - 1)) * p.N;
        s = sp[go]; z = zp[go];
      };
      uint16_t sc0, sc1, sc2;
      uint8_t zz0, zz1, zz2;
      meta_at(0, sc0, zz0); meta_at(1, sc1, zz1); meta_at(2, sc2, zz2);
      for (int st = 0; st < steps; ++st) {
        const int st_in_box = st & 1;
        const uint16_t sc = sc0;
        const uint8_t zz = zz0;
        sc0 = sc1; zz0 = zz1; sc1 = sc2; zz1 = zz2;
        meta_at(st + 3, sc2, zz2);
        if (st_in_box == 0) mbar_wait(pk_full(ps), pph);
        const int chunk = st_in_box * 2 + khalf;       // 16-byte chunk (32 codes) of this row's 64-byte packed line
        const uint4 wv = *reinterpret_cast<const uint4*>(smem_gen + (pk_base - smem_base) + ps * WB_PK_BYTES + pk_row + ((chunk ^ r64) << 4));
        Dq<XT> dq;
        dq.set(sc, (uint32_t)zz);
        const uint4 o0 = dq.cvt(wv.x), o1 = dq.cvt(wv.y), o2 = dq.cvt(wv.z), o3 = dq.cvt(wv.w);
        mbar_wait(b_empty(bs), bph ^ 1u);
        uint8_t* brow = smem_gen + (b_base - smem_base) + bs * WB_B_BYTES + row * 128;
        *reinterpret_cast<uint4*>(brow + (((khalf * 4 + 0) ^ r7) << 4)) = o0;
        *reinterpret_cast<uint4*>(brow + (((khalf * 4 + 1) ^ r7) << 4)) = o1;
        *reinterpret_cast<uint4*>(brow + (((khalf * 4 + 2) ^ r7) << 4)) = o2;
        *reinterpret_cast<uint4*>(brow + (((khalf * 4 + 3) ^ r7) << 4)) = o3;
        fence_proxy_async();                           // generic-proxy stores -> visible to the tensor core's reads
        __syncwarp();
        const bool box_done = st_in_box == 1 || st + 1 == steps;
        if (lane == 0) {
          mbar_arrive(b_full(bs));
          if (box_done) mbar_arrive(pk_empty(ps));
        }
        if (box_done) { if (++ps == WB_PS) { ps = 0; pph ^= 1u; } }
        if (++bs == WB_BS) { bs = 0; bph ^= 1u; }
      }
    }
  } else if (warp >= 4 + WB_DQ_WARPS) {
    // ===================== epilogue: TMEM lane = batch row, columns = weight rows =====================
    const int we = warp & 3;                          // TMEM lane quarter
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
      int n0, k0;
      unit_steps(u, n0, k0);
      const int acc = it & 1;
      mbar_wait_backoff(tfull(acc), (it >> 1) & 1u);
      tc_fence_after();
      const int m = p.m_base + 32 * we + lane;
      const uint32_t t_row = tmem_base + (uint32_t(32 * we) << 16) + acc * WB_TILE_N;
      const bool any_row = p.m_base + 32 * we < p.M;   // warp-uniform: this lane quarter holds real batch rows
      if (any_row) {
        for (int c = 0; c < WB_TILE_N; c += 16) {
          uint32_t v[16];
          tmem_ld_32x16(t_row + c, v);
          tmem_ld_wait();
          if (m < p.M) {
            const int nb = n0 + c;
            if (p.split) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (nb + i < p.N) atomicAdd(p.yacc + (size_t)m * p.N + nb + i, __uint_as_float(v[i]));
            } else if (nb + 16 <= p.N && (p.N & 7) == 0) {
              XT o[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = cvt_out<XT>(__uint_as_float(v[i]));
              uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<XT*>(p.y) + (size_t)m * p.N + nb);
              dst[0] = *reinterpret_cast<const uint4*>(o);
              dst[1] = *reinterpret_cast<const uint4*>(o + 8);
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (nb + i < p.N) reinterpret_cast<XT*>(p.y)[(size_t)m * p.N + nb + i] = cvt_out<XT>(__uint_as_float(v[i]));
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty(acc));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<1>(tmem_base, 512);
}

template <typename XT>
int run_w4a16_wb(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
                 void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  static PerDeviceFlag attr_set;   // function attributes are per device
  static int n_sms = 0;
  if (!attr_set) {
    if (cudaFuncSetAttribute(w4a16_wb_kernel<XT>, cudaFuncAttributeMaxDynamicSharedMemorySize, WB_SMEM) != cudaSuccess)
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

  // k chunk: the split that minimises (waves x boxes per unit + 1 box of fill/drain per unit)
  const int n_tiles = (N + WB_TILE_N - 1) / WB_TILE_N;
  const int boxes = (K + WB_KBOX - 1) / WB_KBOX;
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

  CUtensorMap map_w, map_x;
  int rc = make_tmap_2d_raw(&map_w, qw, CU_TENSOR_MAP_DATA_TYPE_UINT32, (uint64_t)(K / 8), (uint64_t)N, (uint64_t)(K / 8) * 4,
                            16, WB_TILE_N, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc != GQ_OK) return rc;
  rc = make_tmap_2d_raw(&map_x, xs, fmt == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                        (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != GQ_OK) return rc;
  const int split = k_chunks > 1;
  if (split && cudaMemsetAsync(yacc, 0, (size_t)M * N * sizeof(float), st) != cudaSuccess) return GQ_ERR_CUDA;
  for (int m_base = 0; m_base < M; m_base += 128) {
    WbParams p;
    p.scale = scale; p.qzero = qzero; p.y = y; p.yacc = yacc;
    p.M = M; p.N = N; p.K = K; p.G = K / group;
    p.m_base = m_base; p.k_chunk = per * WB_KBOX; p.n_tiles = n_tiles; p.k_chunks = k_chunks; p.split = split;
    p.group_shift = 0;
    while ((1 << p.group_shift) < group) ++p.group_shift;
    const long units = (long)n_tiles * k_chunks;
    const int grid = (int)(units < n_sms ? units : n_sms);
    GQ_LAUNCHED();
    w4a16_wb_kernel<XT><<<grid, WB_THREADS, WB_SMEM, st>>>(map_w, map_x, p);
    if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  }
  if (split) {
    const size_t n = (size_t)M * N;
    GQ_LAUNCHED();
    w4a16_finalize_kernel<XT><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(yacc, reinterpret_cast<XT*>(y), n);
  }
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace

// Shapes the weights-on-N kernel takes: whole 64-k steps, power-of-two groups of at least one 32-k dequant half.
bool w4a16_wb_supports(int M, int N, int K, int group) {
  return M >= 1 && N >= 1 && K >= 64 && (K % 64) == 0 && group >= 32 && (group & (group - 1)) == 0 && (K % group) == 0;
}

int w4a16_wb(const void* x, const uint32_t* qw, const void* scale, const uint8_t* qzero, const int32_t* perm, void* y,
             void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (M <= 0 || N <= 0) return GQ_OK;
  if (!w4a16_wb_supports(M, N, K, group)) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qw) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 255)) return GQ_ERR_ALIGN;
  if (!workspace) return GQ_ERR_ARG;
  if (fmt == 0) return run_w4a16_wb<__half>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
  if (fmt == 1) return run_w4a16_wb<__nv_bfloat16>(x, qw, scale, qzero, perm, y, workspace, M, N, K, group, fmt, st);
  return GQ_ERR_DTYPE;
}

}  // namespace gq
