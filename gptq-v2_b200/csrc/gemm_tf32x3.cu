// fp32-faithful "A^T B" GEMM on the 5th-gen tensor cores:   C += alpha * A^T B
//   A stored (K x M) row-major, B stored (K x N) row-major, C (M x N) row-major, all fp32.
//
// Used for the two K=128-deep updates that dominate everything after the Hessian:
//   K4  lazy-batch trailing update  W[b:, :] -= Hinv[a:b, b:]^T err[a:b, :]   (gptq_py.py:193,
//       gptq_cpp.cpp:81-83 -- an fp32 addmm_ in the reference, TF32 disabled by main.py:12-13)
//   K2  Cholesky trailing update    A22 -= P^T P  (upper tiles only)
// tcgen05 has no fp32 x fp32 kind, and a single TF32 pass (10-bit mantissa) flips quantisation
// codes, so each operand is split on the fly  x = hi + lo,  hi = rna_tf32(x), lo = x - hi,  and
//   A^T B ~= hi^T hi + hi^T lo + lo^T hi        (3 x kind::tf32, fp32 accumulation in TMEM)
// whose relative error (~2^-21) is below what reordering an fp32 sum already causes.
//
// Pipeline per CTA (persistent over 128 x 256 output tiles, or CTA pairs over 256 x 256 tiles: Cfg<2>; 512 threads):
//   warp 0      TMA producer: [16 k][32 cols] fp32 boxes, SWIZZLE_128B_ATOM_32B (both operands MN-major)
//   warps 8-11  split: read the landed tile, write hi back in place and lo next to it
//               (element-wise on identical swizzled addresses, so no layout knowledge needed)
//   warp 1      one thread issues 3 x tcgen05.mma.kind::tf32 per 8-deep k-slice
//   warps 4-7, 12-15  epilogue (32 rows x 128 columns per warp): the k range of a tile leaves tensor memory in chunks
//               (tcgen05.ld) that are summed in REGISTERS in fp32 round-to-nearest (128 accumulators per thread, paid for
//               with setmaxnreg: the control and split warps give registers up); at the end of the tile
//               * alpha -> swizzled smem -> TMA reduce-add (cp.reduce.async.bulk.tensor ... add.f32): C is never read by
//               the SM, the read-modify-write happens in L2, edges are clipped by the tensor map.
#include "gptq_b200_internal.h"
#include "ptx.cuh"

namespace gq {
namespace {

constexpr int TN = 256, BK = 16;
constexpr int SUB = BK * 128;                       // one [16 k][32 fp32] box = 2 KB
constexpr int SPLIT_THREADS = 128;                  // warps 8..11 split the landed tile into hi / lo
constexpr int GEMM_THREADS = 512;                   // 4 control warps, 4 + 4 epilogue warps, 4 split warps
constexpr int EPI_WARPS = 8;
constexpr int EPI_BUF = 32 * 128;                   // one 32 x 32 fp32 staging tile
// CG = 1: one CTA per 128 x 256 output tile.  CG = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) per 256 x 256
// tile: each CTA stages its own 128 rows of A and only HALF of B (128 columns), so the shared-memory traffic per flop
// -- what bounds this kernel -- drops by a third (per 16-deep k-block and CTA: 16 KB landed by TMA, 16 + 16 KB through
// the split stage, 48 KB read by the MMAs = 96 KB instead of 144 KB for the same 768 tensor-core clocks).
template <int CG> struct Cfg {
  static constexpr int TM = 128 * CG;                             // rows of the tile (128 per CTA)
  static constexpr int A_SUBS = 4;                                // 128 rows per CTA
  static constexpr int B_SUBS = (TN / 32) / CG;                   // 8, or this CTA's half: 4
  static constexpr int HI_BYTES = (A_SUBS + B_SUBS) * SUB;        // 24 KB / 16 KB
  static constexpr int STAGE_BYTES = 2 * HI_BYTES;                // hi + lo
  static constexpr int STAGES = (CG == 2) ? 6 : 4;                // 192 KB either way
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_WARPS * EPI_BUF + 1024 + 256;
};

struct GemmTf32Params {
  int count;               // problems in this launch (identical tile grid, individually clipped)
  int tiles_m, tiles_n;    // tile grid of the largest problem
  float alpha;
  int flags;               // TF32_* (gptq_b200_internal.h)
  int kchunk;              // 16-deep k-blocks accumulated in tensor memory before the partial sum leaves through the epilogue
                           // (0 = the whole k range at once)
  Tf32Problem p[TF32_MAX_BATCH];
};

struct TileInfo { int m0, n0, kb0, kb1, prob; bool skip; };

template <int CG>
__device__ __forceinline__ TileInfo decode_tile(const GemmTf32Params& P, int t) {
  constexpr int TM = Cfg<CG>::TM;
  TileInfo ti;
  const int per = P.tiles_m * P.tiles_n;
  ti.prob = t / per;
  const int r = t - ti.prob * per;
  ti.m0 = (r / P.tiles_n) * TM;
  ti.n0 = (r % P.tiles_n) * TN;
  const Tf32Problem& q = P.p[ti.prob];
  ti.skip = ti.m0 >= q.M || ti.n0 >= q.N || ((P.flags & TF32_UPPER_ONLY) && ti.m0 >= ti.n0 + TN);
  // triangular operands shorten the k range (16-deep blocks; the rest of the triangle inside a block is stored as zeros)
  ti.kb0 = (P.flags & TF32_B_LOWER) ? ti.n0 / BK : 0;                                  // B[k][n] = 0 for k < n
  const int kend = (P.flags & TF32_A_UPPER) ? min(q.K, ti.m0 + TM) : q.K;              // A[k][m] = 0 for k > m
  ti.kb1 = kend / BK;
  return ti;
}

__device__ __forceinline__ void tma_reduce_add_2d(const void* tmap, uint32_t src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(reinterpret_cast<uint64_t>(tmap)), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(reinterpret_cast<uint64_t>(tmap)), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ float rna_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

template <int CG>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_a3,
                   const __grid_constant__ CUtensorMap map_b3, const GemmTf32Params p) {
  using C = Cfg<CG>;
  constexpr int TM = C::TM, A_SUBS = C::A_SUBS, B_SUBS = C::B_SUBS, HI_BYTES = C::HI_BYTES, STAGE_BYTES = C::STAGE_BYTES,
                STAGES = C::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t epi_base = smem_base + STAGES * STAGE_BYTES;
  const uint32_t bar_base = epi_base + EPI_WARPS * EPI_BUF;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto ready_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto empty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (3 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (3 * STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (3 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = p.count * p.tiles_m * p.tiles_n;
  // CG = 2: CTA `rank` of the pair owns rows [128 rank, 128 rank + 128) of the tile and stages columns
  // [128 rank, 128 rank + 128) of B; the pair walks the tile list together
  const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;
  const int walker = (CG == 2) ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int n_walkers = (CG == 2) ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int row_off = 128 * rank;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a); tma_prefetch_desc(&map_b); tma_prefetch_desc(&map_c);
    tma_prefetch_desc(&map_a3); tma_prefetch_desc(&map_b3);
    // full: this CTA's own TMA bytes (its split warps wait on it); ready: one arrival per CTA of the pair (after its split warps have
    // synchronised) on the leader's barrier; empty / tfull: the leader's tcgen05.commit, multicast to both CTAs; tempty: 4 epilogue warps per CTA
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(ready_bar(s), CG == 2 ? 2 : SPLIT_THREADS); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), EPI_WARPS * CG); }
    fence_mbar_init();
  }
  if (warp == 2) { tmem_alloc<CG>(tmem_slot, 512); tmem_relinquish<CG>(); }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // everything above is independent of the predecessor kernel; from here on global memory is read and written
  pdl_launch_dependents();
  pdl_wait();
  // register file: 512 threads start with 128 registers each; the control (0-3) and split (8-11) warpgroups keep 80, the two
  // epilogue warpgroups (4-7, 12-15) take 176 for their 128 fp32 accumulators per thread  (128 * (80 + 80 + 176 + 176) = 64 K)
  // (issued at the top of every role branch: ptxas budgets the registers of the code a setmaxnreg dominates)
  const bool is_epi = (warp >= 4 && warp < 8) || warp >= 12;

  if (warp < 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = walker; t < n_tiles; t += n_walkers) {
        const TileInfo ti = decode_tile<CG>(p, t);
        if (ti.skip) continue;
        const Tf32Problem& q = p.p[ti.prob];
        const int m0 = ti.m0 + row_off, n0 = ti.n0 + (CG == 2 ? row_off : 0);
        for (int kb = ti.kb0; kb < ti.kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          mbar_expect_tx(full_bar(stage), HI_BYTES);
          if (p.flags & TF32_COLGROUPS) {
            // one instruction per operand: A_SUBS / B_SUBS consecutive [16 k][32 col] sub-tiles (a box per sub-tile costs the
            // TMA issue slot ~100 cycles each, 8 - 12 of them per 16-deep k-block of ~760 tensor cycles)
            tma_load_3d(sa, &map_a3, full_bar(stage), 0, q.a_row0 + kb * BK, (q.a_col0 + m0) >> 5);
            tma_load_3d(sa + A_SUBS * SUB, &map_b3, full_bar(stage), 0, q.b_row0 + kb * BK, (q.b_col0 + n0) >> 5);
          } else {
#pragma unroll
            for (int i = 0; i < A_SUBS; ++i)
              tma_load_2d(sa + i * SUB, &map_a, full_bar(stage), q.a_col0 + m0 + 32 * i, q.a_row0 + kb * BK);
#pragma unroll
            for (int j = 0; j < B_SUBS; ++j)
              tma_load_2d(sa + (A_SUBS + j) * SUB, &map_b, full_bar(stage), q.b_col0 + n0 + 32 * j, q.b_row0 + kb * BK);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && lane == 0) {                      // CG = 2: the leader issues for the pair
      const uint32_t idesc = umma_idesc(2u /*tf32*/, 1u, 1u, TM, TN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int t = walker; t < n_tiles; t += n_walkers) {
        const TileInfo ti = decode_tile<CG>(p, t);
        if (ti.skip) continue;
        // Deep k ranges leave tensor memory in chunks: the tensor core adds every MMA into the fp32 accumulator with
        // truncation, a bias that grows with the number of accumulations (measured: 2e-5 relative on the 5504-deep
        // products of the triangular inverse at D = 11008); a chunk's partial sum is added to C by the epilogue's
        // reduce-add in L2, which rounds to nearest, while the next chunk accumulates in the other TMEM buffer.
        const int kstep = p.kchunk > 0 ? p.kchunk : (ti.kb1 - ti.kb0 > 0 ? ti.kb1 - ti.kb0 : 1);
        for (int kc = ti.kb0; kc < ti.kb1 || kc == ti.kb0; kc += kstep) {
        const int kce = min(kc + kstep, ti.kb1);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        ++it;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * TN;
        for (int kb = kc; kb < kce; ++kb) {
          mbar_wait(ready_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_hi = smem_base + stage * STAGE_BYTES;
          const uint32_t b_hi = a_hi + A_SUBS * SUB;
          const uint32_t a_lo = a_hi + HI_BYTES;
          const uint32_t b_lo = b_hi + HI_BYTES;
#pragma unroll
          for (int k8 = 0; k8 < BK / 8; ++k8) {
            // MN-major tf32: SWIZZLE_128B_BASE32B, atoms of [4 k][32 fp32]; an MMA (K = 8) spans two
            // atoms (SBO = 512 B) and 32-column groups sit LBO = one sub-block apart; +1024 B per MMA
            const uint64_t ah = umma_smem_desc(a_hi + k8 * 1024, SUB, 512, 1u);
            const uint64_t bh = umma_smem_desc(b_hi + k8 * 1024, SUB, 512, 1u);
            const uint64_t al = umma_smem_desc(a_lo + k8 * 1024, SUB, 512, 1u);
            const uint64_t bl = umma_smem_desc(b_lo + k8 * 1024, SUB, 512, 1u);
            mma_tf32<CG>(d_tmem, al, bh, idesc, ((kb - kc) | k8) ? 1u : 0u);     // small terms first
            mma_tf32<CG>(d_tmem, ah, bl, idesc, 1u);
            mma_tf32<CG>(d_tmem, ah, bh, idesc, 1u);
          }
          if constexpr (CG == 2) tc_commit_cg2_mcast(empty_bar(stage), 0x3); else tc_commit_cg1(empty_bar(stage));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        if constexpr (CG == 2) tc_commit_cg2_mcast(tfull_bar(acc), 0x3); else tc_commit_cg1(tfull_bar(acc));
        }
      }
    }
  } else if (warp >= 8 && warp < 12) {
    // ---- split hi / lo, element-wise over the 24 KB that just landed
    asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
    const int tid = threadIdx.x - 256;                 // 0..SPLIT_THREADS-1
    int stage = 0; uint32_t phase = 0;
    const uint32_t lead_ready0 = (CG == 2) ? mapa(ready_bar(0), 0) : 0u;     // the leader's ready barriers
    for (int t = walker; t < n_tiles; t += n_walkers) {
      const TileInfo ti = decode_tile<CG>(p, t);
      if (ti.skip) continue;
      for (int kb = ti.kb0; kb < ti.kb1; ++kb) {
        mbar_wait(full_bar(stage), phase);
        float4* hi = reinterpret_cast<float4*>(smem_gen + stage * STAGE_BYTES);
        float4* lo = reinterpret_cast<float4*>(smem_gen + stage * STAGE_BYTES + HI_BYTES);
        if (p.flags & TF32_RAW_HI) {
          // the tensor core reads only the upper 19 bits of a kind::tf32 operand, so the landed fp32 tile already IS
          // hi = trunc_tf32(x); only lo = rna_tf32(x - trunc_tf32(x)) is written (one third less shared-memory traffic
          // in this stage: 24 KB read + 24 KB written per k-block instead of 24 + 48)
#pragma unroll
          for (int i = 0; i < HI_BYTES / 16 / SPLIT_THREADS; ++i) {
            const int idx = i * SPLIT_THREADS + tid;
            const float4 v = hi[idx];
            float4 l;
            l.x = rna_tf32(v.x - __uint_as_float(__float_as_uint(v.x) & 0xffffe000u));
            l.y = rna_tf32(v.y - __uint_as_float(__float_as_uint(v.y) & 0xffffe000u));
            l.z = rna_tf32(v.z - __uint_as_float(__float_as_uint(v.z) & 0xffffe000u));
            l.w = rna_tf32(v.w - __uint_as_float(__float_as_uint(v.w) & 0xffffe000u));
            lo[idx] = l;
          }
        } else {
#pragma unroll
          for (int i = 0; i < HI_BYTES / 16 / SPLIT_THREADS; ++i) {
            const int idx = i * SPLIT_THREADS + tid;
            const float4 v = hi[idx];
            float4 h, l;
            h.x = rna_tf32(v.x); h.y = rna_tf32(v.y); h.z = rna_tf32(v.z); h.w = rna_tf32(v.w);
            l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;
            hi[idx] = h;
            lo[idx] = l;
          }
        }
        fence_proxy_async();                            // generic-proxy writes -> visible to the tensor core
        if constexpr (CG == 2) {
          // one remote arrival per CTA and stage (128 cluster-scope arrivals per k-block cost more than the split itself)
          asm volatile("bar.sync 1, 128;" ::: "memory");
          // default semantics (.release.cta) on the leader's barrier, the form CUTLASS' ClusterBarrier::arrive(cta_id)
          // uses: the data this signals is read by THIS SM's tensor core (async proxy, ordered by the fence above), not
          // by the leader's threads; a .release.cluster arrive here costs a cluster-scope fence per k-block and halves
          // the kernel's throughput (measured: 120 vs 247 TFLOP/s at K = 4096)
          if (tid == 0)
            asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" :: "r"(lead_ready0 + 8u * stage) : "memory");
        } else {
          mbar_arrive(ready_bar(stage));
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (is_epi) {
    // ---- epilogue: warp (we, half) owns TMEM lanes [32 we, 32 we + 32) (its hardware lane quarter) and accumulator columns
    // [128 half, 128 half + 128): one output row and 128 columns per thread
    asm volatile("setmaxnreg.inc.sync.aligned.u32 176;");
    const int we = warp & 3, half = warp >= 12 ? 1 : 0;
    const uint32_t stg = epi_base + (we + 4 * half) * EPI_BUF;
    int it = 0;
    for (int t = walker; t < n_tiles; t += n_walkers) {
      const TileInfo ti = decode_tile<CG>(p, t);
      if (ti.skip) continue;
      const Tf32Problem& pq = p.p[ti.prob];
      const int m0 = ti.m0 + row_off, n0 = ti.n0 + 128 * half;
      const int kstep = p.kchunk > 0 ? p.kchunk : (ti.kb1 - ti.kb0 > 0 ? ti.kb1 - ti.kb0 : 1);
      float accv[128];
#pragma unroll
      for (int i = 0; i < 128; ++i) accv[i] = 0.f;
#pragma unroll 1
      for (int kc = ti.kb0; kc < ti.kb1 || kc == ti.kb0; kc += kstep) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1u;
        ++it;
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (uint32_t(32 * we) << 16) + acc * TN + 128 * half;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 32; ++q) accv[32 * c + q] += __uint_as_float(v[q]);   // fp32 round-to-nearest: the chunk sums never meet TMEM's truncating adder
        }
        // this warp has read its part of the accumulator buffer: hand it back to the MMA issuer
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CG == 2) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
          else mbar_arrive(tempty_bar(acc));
        }
      }
      // ---- the tile's sum leaves: * alpha -> swizzled staging tile -> one TMA store / reduce-add per 32 x 32 box
      const bool rows_live = (m0 + 32 * we) < pq.M;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (rows_live && n0 + c * 32 < pq.N) {
          // the copy issued from this staging tile for the previous box must have finished reading it
          if (lane == 0) bulk_wait_read0();
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 o;
            o.x = p.alpha * accv[32 * c + 4 * q + 0]; o.y = p.alpha * accv[32 * c + 4 * q + 1];
            o.z = p.alpha * accv[32 * c + 4 * q + 2]; o.w = p.alpha * accv[32 * c + 4 * q + 3];
            // SWIZZLE_128B staging: row = lane, 16-byte chunk q lands at q ^ (row & 7)
            const uint32_t addr = stg + lane * 128 + ((q ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};"
                         :: "r"(addr), "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w) : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            if (p.flags & TF32_STORE) tma_store_2d(&map_c, stg, pq.c_col0 + n0 + c * 32, pq.c_row0 + m0 + 32 * we);
            else tma_reduce_add_2d(&map_c, stg, pq.c_col0 + n0 + c * 32, pq.c_row0 + m0 + 32 * we);
            bulk_commit();
          }
        }
      }
    }
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, 512);
}

int make_map_f32(CUtensorMap* out, const float* base, int rows, int cols, int ld, int box_cols, int box_rows,
                 int atom32) {
  return make_tmap_2d(out, base, 2, (uint64_t)cols, (uint64_t)rows, (uint64_t)ld * 4, box_cols, box_rows, atom32);
}

}  // namespace

// Tensor maps are over whole matrices; the operands are windows given by coordinates.
int gemm_tf32x3_maps(Tf32Maps* maps, const float* A, int a_rows, int a_cols, int lda, const float* B, int b_rows,
                     int b_cols, int ldb, float* C, int c_rows, int c_cols, int ldc) {
  int rc;
  if ((lda | ldb | ldc) & 3) return GQ_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(C)) & 15)
    return GQ_ERR_ALIGN;
  if ((rc = make_map_f32(&maps->a, A, a_rows, a_cols, lda, 32, BK, 1)) != GQ_OK) return rc;
  if ((rc = make_map_f32(&maps->b, B, b_rows, b_cols, ldb, 32, BK, 1)) != GQ_OK) return rc;
  if ((rc = make_map_f32(&maps->c, C, c_rows, c_cols, ldc, 32, 32, 0)) != GQ_OK) return rc;
  maps->groups_ok = (a_cols % 32 == 0) && (b_cols % 32 == 0);
  if (maps->groups_ok) {
    if ((rc = make_tmap_f32_colgroups(&maps->a3, A, (uint64_t)a_rows, (uint64_t)a_cols, (uint64_t)lda, BK, 4)) != GQ_OK) return rc;
    if ((rc = make_tmap_f32_colgroups(&maps->b3_8, B, (uint64_t)b_rows, (uint64_t)b_cols, (uint64_t)ldb, BK, 8)) != GQ_OK) return rc;
    if ((rc = make_tmap_f32_colgroups(&maps->b3_4, B, (uint64_t)b_rows, (uint64_t)b_cols, (uint64_t)ldb, BK, 4)) != GQ_OK) return rc;
  } else {
    maps->a3 = maps->a; maps->b3_8 = maps->b; maps->b3_4 = maps->b;      // never used: the kernel takes the 2D boxes
  }
  return GQ_OK;
}

template <int CG>
static int launch_gemm(const Tf32Maps& maps, const GemmTf32Params& p, int n_sms, cudaStream_t st) {
  static PerDeviceFlag attr_set;   // function attributes are per device
  if (!attr_set) {
    if (cudaFuncSetAttribute(gemm_tf32x3_kernel<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<CG>::SMEM_BYTES) !=
        cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int tiles = p.count * p.tiles_m * p.tiles_n;
  int walkers = n_sms / CG;
  if (walkers > tiles) walkers = tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CG * walkers);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = Cfg<CG>::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl ? 2 : 1;
  GQ_LAUNCHED();
  return cudaLaunchKernelEx(&cfg, gemm_tf32x3_kernel<CG>, maps.a, maps.b, maps.c, maps.a3, CG == 2 ? maps.b3_4 : maps.b3_8, p) == cudaSuccess
             ? GQ_OK : GQ_ERR_CUDA;
}

int gemm_tf32x3_batch(const Tf32Maps& maps, const Tf32Problem* probs, int count, float alpha, int flags, cudaStream_t st) {
  if (count <= 0) return GQ_OK;
  if (count > TF32_MAX_BATCH) return GQ_ERR_ARG;
  static int n_sms = 0;
  if (n_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sms <= 0) n_sms = 148;
  }
  GemmTf32Params p;
  p.count = 0; p.tiles_m = 0; p.tiles_n = 0; p.alpha = alpha;
  p.flags = flags | (g_use_tf32x3 >= 2 ? TF32_RAW_HI : 0);
  p.kchunk = g_tf32_kchunk;
  bool groups = maps.groups_ok != 0;
  int max_m = 0, max_n = 0;
  for (int i = 0; i < count; ++i) {
    const Tf32Problem& q = probs[i];
    if (q.M <= 0 || q.N <= 0 || q.K <= 0) continue;
    if (q.K % BK != 0) return GQ_ERR_SHAPE;
    // the tensor maps clip at the edge of the WHOLE matrix, not of the window: a store-mode window must be tile-exact
    // in N (32-column boxes) and M (32-row boxes) unless it ends at the matrix edge -- callers guarantee % 32 == 0
    p.p[p.count++] = q;
    groups = groups && ((q.a_col0 | q.b_col0) & 31) == 0;
    max_m = max(max_m, q.M);
    max_n = max(max_n, q.N);
  }
  if (p.count == 0) return GQ_OK;
  if (groups) p.flags |= TF32_COLGROUPS;
  p.tiles_n = (max_n + TN - 1) / TN;
  // CTA pairs (256-row tiles) whenever they need no more waves than single CTAs would: a pair's main loop moves a third
  // less through shared memory per flop, which also shortens the one-tile-per-CTA near updates on the critical path
  const int tiles1 = p.count * ((max_m + 127) / 128) * p.tiles_n;
  const int tiles2 = p.count * ((max_m + 255) / 256) * p.tiles_n;
  const int waves1 = (tiles1 + n_sms - 1) / n_sms, waves2 = (tiles2 + n_sms / 2 - 1) / (n_sms / 2);
  const bool pair = g_use_tf32x3 == 3 && max_m > 128 && waves2 <= waves1;
  if (pair) {
    p.tiles_m = (max_m + 255) / 256;
    return launch_gemm<2>(maps, p, n_sms, st);
  }
  p.tiles_m = (max_m + 127) / 128;
  return launch_gemm<1>(maps, p, n_sms, st);
}

int gemm_tf32x3(const Tf32Maps& maps, int M, int N, int K, int a_row0, int a_col0, int b_row0, int b_col0, int c_row0,
                int c_col0, float alpha, int upper_only, cudaStream_t st) {
  Tf32Problem q{M, N, K, a_col0, a_row0, b_col0, b_row0, c_col0, c_row0};
  return gemm_tf32x3_batch(maps, &q, 1, alpha, upper_only ? TF32_UPPER_ONLY : 0, st);
}

}  // namespace gq
