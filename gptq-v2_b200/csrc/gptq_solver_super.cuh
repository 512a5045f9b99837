#pragma once
// K3 + the near half of K4 in ONE launch per super-block of up to four 128-column blocks.
//
// The per-block solver (gptq_solver_kernel.cuh) leaves the critical path of a linear as a chain of launch pairs:
// solve block b (20 us) -> one-tile-per-CTA near update of the super-block's later columns (18 us, tensor pipe 28 %
// active: set-up and drain, not math) -> solve block b + 1 ...  Here the chain never leaves the SM:
//
//   * a CTA owns 32 NQ weight rows (NQ = 1 .. 4 quarters of 32; lane = row) for the whole super-block;
//   * the column recurrence of a block is the per-block kernel's (same SubSteps / update4 code, same arithmetic and
//     order, so a block is still bit-exact against the oracle's in-block loop, gptq_py.py:186-192);
//   * the errors of every finished 16-column sub-block go from the chain warp's registers straight into TENSOR MEMORY
//     (tcgen05.st: lane = row, 16 columns = 16 k values, once as the raw fp32 = the tf32 "hi" the MMA truncates to, once
//     as lo = rna_tf32(e - trunc_tf32(e))) and become the A operand of tcgen05.mma.kind::tf32 in TS mode;
//   * B = U[block rows, later columns of the super-block] streams through a TMA ring in [8 k][32 col] boxes
//     (SWIZZLE_128B_ATOM_32B, MN-major) and is split hi / lo element-wise by two warps, exactly as gemm_tf32x3.cu does;
//   * the products E^T-free:  acc[rows][later columns] += E[rows][k] U[k][later columns]  (3 x TF32: lo*hi + hi*lo +
//     hi*hi) accumulate in tensor memory, one 128-column fp32 region per later block, WHILE the chain warp is already
//     in its next sub-block: when block b ends only the MMAs of its last 16 columns are outstanding;
//   * the next block starts from  W_global - acc  (tcgen05.ld by the warps that own those lanes), which replaces the
//     near-update GEMM launches; everything behind the super-block gets one rank-(128 nblk) update from
//     gemm_tf32x3_kernel as before (gptq_py.py:193, summed per super-block).
//
// Warps: 0..NQ-1 chain (one per SM sub-partition), NQ..15 helpers (look-ahead rank-16 updates inside the block and the
// coalesced write-out), 16 TMA producer (+ TMEM allocation), 17-18 hi/lo split, 19 MMA issuer.
#include <cuda_fp16.h>
#include "gptq_b200_internal.h"
#include "gptq_solver_kernel.cuh"
#include "ptx.cuh"

namespace gq {

constexpr int SUP_NB = 128;
constexpr int SUP_COMPUTE_WARPS = 16;
constexpr int SUP_COMPUTE_THREADS = SUP_COMPUTE_WARPS * 32;
constexpr int SUP_THREADS = SUP_COMPUTE_THREADS + 4 * 32;
constexpr int SUP_SPLIT_THREADS = 64;
constexpr int SUP_SUBTILE = 8 * 128;                    // one [8 k][32 fp32] box
constexpr int SUP_STAGE_HI = 12 * SUP_SUBTILE;          // up to 384 later columns per 8-deep k-slice
constexpr int SUP_STAGE_BYTES = 2 * SUP_STAGE_HI;       // hi + lo
constexpr int SUP_TMEM_A = 3 * 128;                     // A operand buffers behind the three accumulator regions
constexpr int SUP_UPK_BYTES = UPK_FLOATS<SUP_NB> * 4;

template <int MODE, int NQ> struct SupCfg {
  // per quarter (32 rows): ws[128][32] | es, qs (, xs) [2][16][32] | mm[2][4][32] | sz[2][32]
  static constexpr int Q_FLOATS = SUP_NB * 32 + (MODE == 1 ? 3 : 2) * 2 * SB * 32 + 2 * 4 * 32 + 2 * 32;
  // the diagonal block of U is double-buffered (block b + 1's copy lands while block b is solved) wherever that still
  // leaves two ring stages; only (MODE 1, 128 rows per CTA) loads it at the block boundary instead
  static constexpr int NUPK = (227 * 1024 - (2 * UPK_FLOATS<SUP_NB> + 2 * SUP_NB + NQ * Q_FLOATS) * 4 - 1024 - 256 >= 2 * SUP_STAGE_BYTES) ? 2 : 1;
  static constexpr int FIXED_FLOATS = NUPK * UPK_FLOATS<SUP_NB> + 2 * SUP_NB + NQ * Q_FLOATS;
  static constexpr int FIXED_BYTES = FIXED_FLOATS * 4;
  static constexpr int AVAIL = 227 * 1024 - FIXED_BYTES - 1024 - 256;
  static constexpr int STAGES = AVAIL / SUP_STAGE_BYTES > 6 ? 6 : AVAIL / SUP_STAGE_BYTES;
  static_assert(STAGES >= 2, "the B ring needs one sub-block (two k-slices) in flight");
  static constexpr int SMEM_BYTES = 1024 + STAGES * SUP_STAGE_BYTES + FIXED_BYTES + 256;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory");
  static_assert(SMEM_BYTES > 114 * 1024, "one CTA per SM (it allocates all 512 TMEM columns)");
};

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
               :: "r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ float sup_rna_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float sup_lo(float x) {            // rna_tf32(x - trunc_tf32(x))
  return sup_rna_tf32(x - __uint_as_float(__float_as_uint(x) & 0xffffe000u));
}

template <int MODE, int NQ, typename CodeT>
__global__ void __launch_bounds__(SUP_THREADS, 1)
solve_super_kernel(const __grid_constant__ SuperMaps maps_u, const SuperParams p) {
  using Cfg = SupCfg<MODE, NQ>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t sup_smem[];
  const uint32_t ring = (smem_u32(sup_smem) + 1023u) & ~1023u;
  uint8_t* ring_gen = sup_smem + (ring - smem_u32(sup_smem));
  constexpr int NUPK = Cfg::NUPK;
  float* Upk0 = reinterpret_cast<float*>(ring_gen + STAGES * SUP_STAGE_BYTES);   // [NUPK] packed diagonal blocks of U
  float* rdg = Upk0 + NUPK * UPK_FLOATS<SUP_NB>;             // [2][128] reciprocals of diag(U), double-buffered per block
  float* qbase = rdg + 2 * SUP_NB;
  const uint32_t bar0 = smem_u32(qbase + NQ * Cfg::Q_FLOATS);
  auto bfull = [&](int s) { return bar0 + 8u * s; };
  auto bready = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto bempty = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto aready = [&](int i) { return bar0 + 8u * (3 * STAGES + i); };
  auto aempty = [&](int i) { return bar0 + 8u * (3 * STAGES + 2 + i); };
  auto accfull = [&](int r) { return bar0 + 8u * (3 * STAGES + 4 + r); };
  auto ubar = [&](int i) { return bar0 + 8u * (3 * STAGES + 7 + i); };
  const uint32_t ufree = bar0 + 8u * (3 * STAGES + 9);       // one arrival per block boundary: the previous block's U buffer is free
  const uint32_t tmem_slot = bar0 + 8u * (3 * STAGES + 10);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(
      reinterpret_cast<uint8_t*>(qbase + NQ * Cfg::Q_FLOATS) + 8u * (3 * STAGES + 10));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int R = p.R;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&maps_u.m[0]); tma_prefetch_desc(&maps_u.m[1]); tma_prefetch_desc(&maps_u.m[2]);
    for (int g = 0; g < 8; ++g) tma_prefetch_desc(&maps_u.upk[g]);
    for (int s = 0; s < STAGES; ++s) { mbar_init(bfull(s), 1); mbar_init(bready(s), SUP_SPLIT_THREADS); mbar_init(bempty(s), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(aready(i), NQ); mbar_init(aempty(i), 1); }
    for (int r = 0; r < 3; ++r) mbar_init(accfull(r), 1);
    mbar_init(ubar(0), 1); mbar_init(ubar(1), 1); mbar_init(ufree, 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  if (warp == 16) { tmem_alloc<1>(tmem_slot, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_launch_dependents();       // the far update behind this super-block may set up while the chain runs
  pdl_wait();                    // the previous far update has landed in W before anything below reads global memory

  if (warp == 16) {
    // ------------------------------------------------------------ TMA producer: U[block rows, later columns] through the ring
    // (lane 0) and the packed diagonal block of U of every block (all lanes, one row copy each: 128 bulk copies take the
    // TMA unit ~7000 cycles, so they are issued here -- not by a warp the chain meets at a barrier -- in batches of 32
    // rows between ring stages, one block ahead where the buffer is double)
    {
      auto upk_begin = [&](int t) {
        if (lane == 0) mbar_expect_tx(ubar(t & 1), SUP_UPK_BYTES);
        __syncwarp();
      };
      auto upk_group = [&](int t, int g) {                       // rows [16 g, 16 g + 16) of block t's diagonal block
        const int cb = p.c0 + SUP_NB * t;
        float* Upk = Upk0 + (t % NUPK) * UPK_FLOATS<SUP_NB>;
        if (lane == 0) tma_load_2d(smem_u32(Upk + upk_base<SUP_NB>(g)), &maps_u.upk[g], ubar(t & 1), cb + 16 * g, cb + 16 * g);
      };
      auto wait_free = [&](int bnd) {                            // boundary `bnd` passed: everyone is done with block bnd - 1
        if (lane == 0) mbar_wait(ufree, bnd & 1u);
        __syncwarp();
      };
      upk_begin(0);
      for (int g = 0; g < SUP_NB / SB; ++g) upk_group(0, g);
      int stage = 0; uint32_t phase = 0;
      for (int b = 0; b + 1 < p.nblk; ++b) {
        const int nsub = 4 * (p.nblk - 1 - b);
        const int col = p.c0 + SUP_NB * (b + 1), row = p.c0 + SUP_NB * b;
        const int t = NUPK == 2 ? b + 1 : b;                     // the diagonal block fetched during this block's ring loop
        if (NUPK == 1 && t >= 1) {                               // single buffer: needed at once
          wait_free(t);
          upk_begin(t);
          for (int g = 0; g < SUP_NB / SB; ++g) upk_group(t, g);
        }
        for (int k8 = 0; k8 < SUP_NB / 8; ++k8) {
          if (lane == 0) {
            mbar_wait(bempty(stage), phase ^ 1u);
            mbar_expect_tx(bfull(stage), nsub * SUP_SUBTILE);
            const uint32_t dst = ring + stage * SUP_STAGE_BYTES;
            // one instruction per k-slice: nsub consecutive [8 k][32 col] sub-tiles (a box per sub-tile kept the TMA issue
            // slot busy for ~3000 cycles per sub-block: the tensor core fell 8000 cycles behind the chain by the end of a block)
            tma_load_3d(dst, &maps_u.m[nsub / 4 - 1], bfull(stage), 0, row + 8 * k8, col / 32);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          if (NUPK == 2 && k8 >= 3 && k8 <= 9 && (k8 & 1)) {     // after k-slices 3, 5, 7, 9: two 16-row groups each
            if (k8 == 3) { wait_free(b); upk_begin(t); }
            upk_group(t, k8 - 3);
            upk_group(t, k8 - 2);
          }
        }
      }
      if (NUPK == 1 && p.nblk >= 2) {                            // the last block has no ring loop of its own
        const int t = p.nblk - 1;
        wait_free(t);
        upk_begin(t);
        for (int g = 0; g < SUP_NB / SB; ++g) upk_group(t, g);
      }
    }
  } else if (warp == 17 || warp == 18) {
    // ------------------------------------------------------------ lo = rna_tf32(x - trunc_tf32(x)) next to the landed tile
    const int tid = threadIdx.x - 17 * 32;
    int stage = 0; uint32_t phase = 0;
    for (int b = 0; b + 1 < p.nblk; ++b) {
      const int n4 = 4 * (p.nblk - 1 - b) * (SUP_SUBTILE / 16);
      for (int k8 = 0; k8 < SUP_NB / 8; ++k8) {
        mbar_wait(bfull(stage), phase);
        const float4* hi = reinterpret_cast<const float4*>(ring_gen + stage * SUP_STAGE_BYTES);
        float4* lo = reinterpret_cast<float4*>(ring_gen + stage * SUP_STAGE_BYTES + SUP_STAGE_HI);
        // n4 = 256 float4 per accumulator region: 4, 8 or 12 per thread, four loads in flight at a time (with one at a time
        // the two split warps only just kept pace with a three-region block and the tensor core waited for them)
#pragma unroll 1
        for (int i = tid; i < n4; i += 4 * SUP_SPLIT_THREADS) {
          float4 v[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = hi[i + k * SUP_SPLIT_THREADS];
#pragma unroll
          for (int k = 0; k < 4; ++k)
            lo[i + k * SUP_SPLIT_THREADS] = make_float4(sup_lo(v[k].x), sup_lo(v[k].y), sup_lo(v[k].z), sup_lo(v[k].w));
        }
        fence_proxy_async();
        mbar_arrive(bready(stage));
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 19) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc128 = umma_idesc(2u, 0u, 1u, 128, 128), idesc256 = umma_idesc(2u, 0u, 1u, 128, 256);
      int stage = 0; uint32_t phase = 0;
      int u = 0;
      long long* mdbg = (p.dbg && blockIdx.x == 0) ? p.dbg + 256 * 4 : nullptr;   // profiling: (aready, bready, issued) per sub-block
      int mdi = 0;
      auto mstamp = [&]() { if (mdbg && mdi < 256) mdbg[mdi++] = clock64(); };
      for (int b = 0; b + 1 < p.nblk; ++b) {
        const int nreg = p.nblk - 1 - b;                       // accumulator regions this block feeds
        for (int sb = 0; sb < SUP_NB / SB; ++sb, ++u) {
          const int abuf = u & 1;
          mbar_wait(aready(abuf), (u >> 1) & 1u);
          mstamp();
          const int st0 = stage; const uint32_t ph0 = phase;
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          const int st1 = stage; const uint32_t ph1 = phase;
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          mbar_wait(bready(st0), ph0);
          mbar_wait(bready(st1), ph1);
          mstamp();
          tc_fence_after();
          const uint32_t a_hi = tmem_base + SUP_TMEM_A + abuf * 32, a_lo = a_hi + 16;
          const uint32_t first = (b == 0 && sb == 0) ? 0u : 1u;
          const uint32_t idesc384a = idesc256;                 // three regions = one N = 256 MMA + one N = 128 MMA
          if (sb == SUP_NB / SB - 1) {
            // last sub-block: the next block's 128 columns first (the chain waits for them), then the rest
#pragma unroll
            for (int k8 = 0; k8 < 2; ++k8) {
              const uint32_t bh = ring + (k8 ? st1 : st0) * SUP_STAGE_BYTES, bl = bh + SUP_STAGE_HI;
              // MN-major tf32, SWIZZLE_128B_BASE32B: atoms of [4 k][32 fp32] (SBO = 512 B), 32-column groups one box apart
              const uint64_t dh = umma_smem_desc(bh, SUP_SUBTILE, 512, 1u), dl = umma_smem_desc(bl, SUP_SUBTILE, 512, 1u);
              const uint32_t d = tmem_base + SUP_NB * b;
              mma_tf32_ts(d, a_lo + 8 * k8, dh, idesc128, k8 ? 1u : first);      // small terms first
              mma_tf32_ts(d, a_hi + 8 * k8, dl, idesc128, 1u);
              mma_tf32_ts(d, a_hi + 8 * k8, dh, idesc128, 1u);
            }
            tc_commit_cg1(accfull(b));
            if (nreg >= 2) {
              const uint32_t idesc = nreg == 3 ? idesc256 : idesc128;
#pragma unroll
              for (int k8 = 0; k8 < 2; ++k8) {
                const uint32_t bh = ring + (k8 ? st1 : st0) * SUP_STAGE_BYTES + 4 * SUP_SUBTILE, bl = bh + SUP_STAGE_HI;
                const uint64_t dh = umma_smem_desc(bh, SUP_SUBTILE, 512, 1u), dl = umma_smem_desc(bl, SUP_SUBTILE, 512, 1u);
                const uint32_t d = tmem_base + SUP_NB * (b + 1);
                mma_tf32_ts(d, a_lo + 8 * k8, dh, idesc, k8 ? 1u : first);
                mma_tf32_ts(d, a_hi + 8 * k8, dl, idesc, 1u);
                mma_tf32_ts(d, a_hi + 8 * k8, dh, idesc, 1u);
              }
            }
          } else {
            // every other sub-block: as few MMAs as the N <= 256 limit allows (an MMA holds the tensor core ~172 cycles
            // whatever its N): regions [b, b + 2) in one N = 256 (or N = 128) MMA, a third region in a second one
#pragma unroll
            for (int k8 = 0; k8 < 2; ++k8) {
              const uint32_t bh = ring + (k8 ? st1 : st0) * SUP_STAGE_BYTES, bl = bh + SUP_STAGE_HI;
              const uint64_t dh = umma_smem_desc(bh, SUP_SUBTILE, 512, 1u), dl = umma_smem_desc(bl, SUP_SUBTILE, 512, 1u);
              const uint32_t d = tmem_base + SUP_NB * b;
              const uint32_t idesc = nreg >= 2 ? idesc384a : idesc128;
              mma_tf32_ts(d, a_lo + 8 * k8, dh, idesc, k8 ? 1u : first);
              mma_tf32_ts(d, a_hi + 8 * k8, dl, idesc, 1u);
              mma_tf32_ts(d, a_hi + 8 * k8, dh, idesc, 1u);
              if (nreg == 3) {
                const uint64_t dh2 = umma_smem_desc(bh + 8 * SUP_SUBTILE, SUP_SUBTILE, 512, 1u);
                const uint64_t dl2 = umma_smem_desc(bl + 8 * SUP_SUBTILE, SUP_SUBTILE, 512, 1u);
                const uint32_t d2 = tmem_base + SUP_NB * (b + 2);
                mma_tf32_ts(d2, a_lo + 8 * k8, dh2, idesc128, k8 ? 1u : first);
                mma_tf32_ts(d2, a_hi + 8 * k8, dl2, idesc128, 1u);
                mma_tf32_ts(d2, a_hi + 8 * k8, dh2, idesc128, 1u);
              }
            }
          }
          tc_commit_cg1(bempty(st0));
          tc_commit_cg1(bempty(st1));
          tc_commit_cg1(aempty(abuf));
          mstamp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------ chain + helper warps
    // Roles among the 16 compute warps (quarter q = 32 rows = TMEM lanes 32 q ..):
    //   chain q          warp q (q < NQ): the column recurrence; after publishing a sub-block's errors it also hands them
    //                    to the tensor core (tcgen05.st) while the look-ahead warps update the next 16 columns
    //   stagers of q     warps q + 4, q + 8, q + 12 (the other warps that may touch quarter q's TMEM lanes): at a block
    //                    boundary they load their 48 / 40 / 40 columns of the block's weights (in flight while the previous
    //                    block's last MMAs finish) and turn them into W - acc
    //   look-ahead       the rank-16 update of the next sub-block's 16 columns, four 4-column items per quarter: on four
    //                    non-chain warps of different SM sub-partitions (NQ <= 2), or on the quarter's chain warp + its three
    //                    stagers (NQ = 4, where there are no other warps)
    //   far / write-out  warps that do not share an SM sub-partition with a chain warp (all non-chain warps when NQ = 4):
    //                    the rest of the block's columns and the global stores, while the chain runs its next sub-block
    const int row0 = blockIdx.x * 32 * NQ;
    const int my_q = warp & 3;
    auto Qp = [&](int q) { return qbase + q * Cfg::Q_FLOATS; };
    constexpr int OFF_ES = SUP_NB * 32, OFF_QS = OFF_ES + 2 * SB * 32, OFF_XS = OFF_QS + 2 * SB * 32;
    constexpr int OFF_MM = (MODE == 1 ? OFF_XS + 2 * SB * 32 : OFF_XS), OFF_SZ = OFF_MM + 2 * 4 * 32;
    CodeT* codes = reinterpret_cast<CodeT*>(p.codes_t);
    const bool is_chain = warp < NQ;
    const bool is_stager = !is_chain && my_q < NQ && warp >= 4;
    const int part = (warp >> 2) - 1;                            // stagers: 0, 1, 2
    const int part_c0 = part == 0 ? 0 : part == 1 ? 48 : 88;
    const int part_w = part == 0 ? 48 : 40;
    // look-ahead item of this warp: (quarter la_q, column group la_g) or none (la_q < 0)
    int la_q = -1, la_g = 0;
    constexpr bool DENSE = NQ >= 2;          // two or more chain warps: a quarter's look-ahead stays on its own SM sub-partition (NQ = 1: on the three others, 0.89 vs 0.98 ms)
    if (DENSE) { if (my_q < NQ) { la_q = my_q; la_g = warp >> 2; } }
    else if (warp >= NQ && warp < NQ + 4 * NQ) { la_q = (warp - NQ) >> 2; la_g = (warp - NQ) & 3; }
    // far-update / write-out helpers
    constexpr int NH = DENSE ? 16 - NQ : 4 * (4 - NQ);
    const bool is_far = DENSE ? warp >= NQ : my_q >= NQ;
    const int far_i = DENSE ? warp - NQ : (warp >> 2) * (4 - NQ) + (my_q - NQ);
    // profiling hook: time stamps of CTA 0's chain warp
    // (reporters: lane 0 of warps 0, 1, 4 and 15 of CTA 0, 256 slots each)
    const int rep = warp == 0 ? 0 : warp == 1 ? 1 : warp == 4 ? 2 : warp == 15 ? 3 : -1;
    long long* dbg = (p.dbg && blockIdx.x == 0 && lane == 0 && rep >= 0) ? p.dbg + 256 * rep : nullptr;
    int di = 0;
    auto stamp = [&]() { if (dbg && di < 256) dbg[di++] = clock64(); };

    // reciprocals of diag(U) of block b by threads 128..255: the load is issued at one point and consumed sub-blocks
    // later (a strided global load per thread; its latency must not sit in front of a barrier the chain waits behind)
    float rdg_u = 1.f;
    auto issue_rdg = [&](int b) {
      const int cb = p.c0 + SUP_NB * b, t = threadIdx.x - 128;
      if (t >= 0 && t < SUP_NB)
        asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(rdg_u) : "l"(p.U + (size_t)(cb + t) * p.C + cb + t));
    };
    auto finish_rdg = [&](int b) {
      const int t = threadIdx.x - 128;
      if (t >= 0 && t < SUP_NB) rdg[(b & 1) * SUP_NB + t] = __frcp_rn(rdg_u);
    };
    // stagers: their columns of block b's weights, global -> registers (issued early, consumed at the block boundary)
    float wv[48];
    auto prefetch_w = [&](int b) {
      const int cb = p.c0 + SUP_NB * b;
      const int r = row0 + my_q * 32 + lane;
      const bool act = r < R;
#pragma unroll
      for (int i = 0; i < 48; ++i) wv[i] = (act && i < part_w) ? p.Wt[(size_t)(cb + part_c0 + i) * R + r] : 0.f;
    };

    issue_rdg(0);
    finish_rdg(0);

#pragma unroll 1
    for (int b = 0; b < p.nblk; ++b) {
      const int cb = p.c0 + SUP_NB * b;
      const float* Upk = Upk0 + (b % NUPK) * UPK_FLOATS<SUP_NB>;
      const float* rdg_b = rdg + (b & 1) * SUP_NB;
      // ---- block boundary.  Everyone is done with the previous block's ws / es (and its Upk when single-buffered).
      stamp();
      named_bar_sync(1, SUP_COMPUTE_THREADS);
      stamp();
      if (threadIdx.x == 0) mbar_arrive(ufree);                 // the producer may overwrite the previous block's U buffer
      if (is_stager) {
        // W_global - (updates of the super-block's earlier blocks, accumulated in tensor memory); the global loads are in
        // flight while the last MMAs of the previous block finish
        prefetch_w(b);
        stamp();
        float* ws_q = Qp(my_q);
        if (b > 0) {
          mbar_wait(accfull(b - 1), 0);
          stamp();
          tc_fence_after();
          const uint32_t t = tmem_base + (uint32_t(32 * my_q) << 16) + SUP_NB * (b - 1) + part_c0;
#pragma unroll
          for (int h = 0; h < 3; ++h) {                          // the third load of a 40-column part reads 8 columns too many
            uint32_t a[16];
            tmem_ld_32x16(t + 16 * h, a);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) wv[16 * h + i] = __fsub_rn(wv[16 * h + i], __uint_as_float(a[i]));
          }
        }
        else stamp();
        stamp();
        float mx = -INFINITY, mn = INFINITY;
#pragma unroll
        for (int i = 0; i < 48; ++i) {
          if (i < part_w) {
            ws_q[(part_c0 + i) * 32 + lane] = wv[i];
            mx = fmaxf(mx, wv[i]); mn = fminf(mn, wv[i]);
          }
        }
        ws_q[OFF_MM + part * 32 + lane] = mx;
        ws_q[OFF_MM + (4 + part) * 32 + lane] = mn;
      } else { stamp(); stamp(); stamp(); }
      stamp();
      named_bar_sync(1, SUP_COMPUTE_THREADS);                    // the block's weights and min / max partials are in place
      // the NEXT block's diagonal block of U and reciprocals start travelling now
      if (b + 1 < p.nblk) {
        issue_rdg(b + 1);
      }
      const float maxq = b == 0 ? p.maxq[0] : b == 1 ? p.maxq[1] : b == 2 ? p.maxq[2] : p.maxq[3];   // no dynamic indexing of the parameter space
      float s = 1.f, z = 0.f;
      if (is_chain) {
        float* ws_q = Qp(warp);
        const int r = row0 + warp * 32 + lane;
        const bool act = r < R;
        float* sc = p.scale + (size_t)b * p.param_stride;
        float* zp = p.qzero + (size_t)b * p.param_stride;
        if (MODE == 0 && !p.static_params) {
          float mx = ws_q[OFF_MM + lane], mn = ws_q[OFF_MM + 4 * 32 + lane];
#pragma unroll
          for (int i = 1; i < 3; ++i) {
            mx = fmaxf(mx, ws_q[OFF_MM + i * 32 + lane]);
            mn = fminf(mn, ws_q[OFF_MM + (4 + i) * 32 + lane]);
          }
          const float xmax = fmaxf(mx, 0.f);                     // .relu()           (quant.py:114)
          const float xmin = -fmaxf(-mn, 0.f);                   // -(-min).relu()    (quant.py:115)
          s = __fadd_rn(__fdiv_rn(__fsub_rn(xmax, xmin), maxq), 1e-12f);
          z = rintf(__fdiv_rn(-xmin, s));
          if (act) { sc[r] = s; zp[r] = z; }
        } else if (act) {
          s = sc[r]; z = zp[r];
        }
        ws_q[OFF_SZ + lane] = s; ws_q[OFF_SZ + 32 + lane] = z;   // read by the write-out warps behind the next barrier
      }
      const float rs = __frcp_rn(s);
      __half hs = __float2half_rn(0.f), hz = hs;
      if (MODE == 1) { hs = __float2half_rn(s); hz = __float2half_rn(z); }
      mbar_wait(ubar(b & 1), (b >> 1) & 1u);                     // every warp: the diagonal block of U has landed
      stamp();

      // write-out of one finished column of one quarter (coalesced 128-byte rows)
      auto store_col = [&](int sb, int rq, int j) {
        const int r = row0 + rq * 32 + lane;
        if (r >= R) return;
        const float* q_ = Qp(rq);
        const int o_s = (sb & 1) * SB * 32 + j * 32 + lane;
        const size_t o = (size_t)(cb + sb * SB + j) * R + r;
        const float qv = q_[OFF_QS + o_s];
        const float ss = q_[OFF_SZ + lane], zz = q_[OFF_SZ + 32 + lane];
        if (MODE == 1) p.Wt[o] = q_[OFF_XS + o_s];
        if (p.quant_t) p.quant_t[o] = dequant_code<MODE>(qv, ss, zz, __float2half_rn(ss), __float2half_rn(zz));
        if (codes) codes[o] = static_cast<CodeT>(qv);
        p.err_t[o] = q_[OFF_ES + o_s];
      };

#pragma unroll 1
      for (int sb = 0; sb < SUP_NB / SB; ++sb) {
        const int col0 = sb * SB;
        const int ld = SUP_NB - col0;
        const float* Usb = Upk + upk_base<SUP_NB>(sb);
        const int par = (sb & 1) * SB * 32;
        const bool last_sb = sb + 1 == SUP_NB / SB;
        stamp();
        if (is_chain) {
          float* ws_q = Qp(warp);
          float w[SB];
#pragma unroll
          for (int j = 0; j < SB; ++j) w[j] = ws_q[(col0 + j) * 32 + lane];
          SubSteps<MODE, 0>::run(w, Usb, ld, rdg_b + col0, ws_q + OFF_ES + par, ws_q + OFF_QS + par, ws_q + OFF_XS + par, s, rs,
                                 z, maxq, hs, hz, lane);
        }
        if (sb == 4 && b + 1 < p.nblk) finish_rdg(b + 1);        // issued at the block's start: long landed
        stamp();
        stamp();
        named_bar_sync(1, SUP_COMPUTE_THREADS);                  // errors / codes of sub-block sb published
        stamp();
        if (!last_sb && la_q >= 0) {
          // the next sub-block's 16 columns first: the recurrence waits only for these
          float* ws_q = Qp(la_q);
          float e[SB];
#pragma unroll
          for (int j = 0; j < SB; ++j) e[j] = ws_q[OFF_ES + par + j * 32 + lane];
          update4<MODE>(ws_q, e, Usb, ld, col0, col0 + SB + 4 * la_g, lane);
        }
        if (is_chain && b + 1 < p.nblk) {
          // the 16 errors of this sub-block -> tensor memory (A operand of the near update), off the look-ahead's path
          float* ws_q = Qp(warp);
          const int u = b * (SUP_NB / SB) + sb, abuf = u & 1;
          if (u >= 2) mbar_wait(aempty(abuf), ((u >> 1) - 1) & 1u);       // the MMAs that read this buffer are done
          tc_fence_after();
          uint32_t hi[SB], lo[SB];
#pragma unroll
          for (int j = 0; j < SB; ++j) {
            const float e = ws_q[OFF_ES + par + j * 32 + lane];
            hi[j] = __float_as_uint(e);
            lo[j] = __float_as_uint(sup_lo(e));
          }
          const uint32_t t = tmem_base + (uint32_t(32 * warp) << 16) + SUP_TMEM_A + abuf * 32;
          tmem_st_32x16(t, hi);
          tmem_st_32x16(t + 16, lo);
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(aready(abuf));
        }
        if (last_sb) break;
        stamp();
        named_bar_sync(1, SUP_COMPUTE_THREADS);
        if (is_far) {
          // write-out + far columns of the block while the chain warps run the next sub-block
          for (int it = far_i; it < SB * NQ; it += NH) store_col(sb, it / SB, it % SB);
          const int ng = (SUP_NB - col0 - 2 * SB) / 4;
#pragma unroll 1
          for (int it = far_i; it < ng * NQ; it += NH) {
            const int rq = it / ng, k = col0 + 2 * SB + 4 * (it - rq * ng);
            float* ws_q = Qp(rq);
            float e[SB];
#pragma unroll
            for (int j = 0; j < SB; ++j) e[j] = ws_q[OFF_ES + par + j * 32 + lane];
            update4<MODE>(ws_q, e, Usb, ld, col0, k, lane);
          }
        }
      }
      // last sub-block of the block
      if (is_far)
        for (int it = far_i; it < SB * NQ; it += NH) store_col(SUP_NB / SB - 1, it / SB, it % SB);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 16) tmem_dealloc<1>(tmem_base, 512);
}

template <int MODE, int NQ, typename CodeT>
int launch_super(const SuperMaps& maps_u, const SuperParams& p, cudaStream_t st) {
  using Cfg = SupCfg<MODE, NQ>;
  static PerDeviceFlag attr_set;   // function attributes are per device
  if (!attr_set) {
    if (cudaFuncSetAttribute(solve_super_kernel<MODE, NQ, CodeT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Cfg::SMEM_BYTES) != cudaSuccess)
      return GQ_ERR_CUDA;
    attr_set = true;
  }
  const int rows = 32 * NQ;
  GQ_LAUNCHED();
  return launch_chain(solve_super_kernel<MODE, NQ, CodeT>, dim3((p.R + rows - 1) / rows), dim3(SUP_THREADS), (size_t)Cfg::SMEM_BYTES, st,
                      g_pdl != 0, maps_u, p) == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
