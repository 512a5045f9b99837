// K6, batch 9..16 (and batch 3..8 when K is too long for w4a16_smallm.cu's resident activation):
//   y (M, N) = x (M, K) . W_hat^T   with W_hat = (q - z) s  from packed int4, one pass over the weights.
//
// Same arithmetic as w4a16_smallm.cu (the packed word becomes the A fragment of a warp-level mma.m16n8k16 with the
// nibble OR-ed into a constant's mantissa; offset and zero point leave per 128-column chunk through the chunk's
// activation sum; scale applied in fp32), with two changes that lift its batch limit:
//
//   * NBK = 1 or 2 blocks of 8 batch columns per weight fragment: the unpack is paid once, the second block costs two
//     MMAs and one 16-byte shared-memory load per packed word pair.  Measured (profiles/r02_v38_*): batch 16 on
//     28672 x 8192 in 58 us against 85 us on the tcgen05 path, whose 128-row MMA floor streams 1.6 TB/s of packed weights
//     at best (w4a16_mma.cu), and 75 us for cuBLAS on the dense weights.  Four blocks (batch 32) were tried and are NOT
//     kept: ncu shows no saturated pipe but ALU + tensor + LSU issue adding up (issue slots 45 % busy with 16 or with 32
//     warps per SM), so every further block costs its full 250 clk per unit and the tcgen05 path wins from batch 17 on.
//   * the activation does not have to fit in shared memory: it is streamed in WINDOWS of `wc` chunks (all batch columns,
//     fragment order, from the pre-gathered copy w4a16_smallm_xprep_kernel leaves in the workspace), double-buffered
//     with cp.async.  A CTA (one per SM) owns whole 16-row tiles; in every round (= window) its (tile, chunk) units are
//     dealt to the 16 warps as equal contiguous ranges, tile-major, so a warp keeps its fp32 sums in registers while it
//     stays in a tile and adds them into the tile's shared-memory sums when it leaves it.  One barrier per round; the
//     weight ring (MM_PF units of 1 KB per warp in flight) runs across the rounds without draining.
//
// Needs group % 128 == 0, K % 128 == 0, N % 16 == 0, M <= 16.
// Reference semantics: reconstruct_nn_linear (quant.py:262-279) followed by nn.Linear on the same input.
#include "w4a16_frag.cuh"

namespace gq {
namespace {

constexpr int MM_MAXM = 16;
constexpr int MM_PF = 6;                     // units in flight per warp (register ring)

struct MidMParams {
  const uint32_t* qw;       // (N, K/8)
  const void* scale_t;      // (G, N)
  const uint8_t* qzero_t;   // (G, N)
  const void* xs_pre;       // (M, K) activation in fragment order (w4a16_smallm_xprep_kernel)
  const float* s_pre;       // (M, K/128) chunk sums
  void* y;                  // (M, N) 16-bit
  int M, N, K, group;
  int tiles_per_cta;        // row tiles (16 rows) owned by one CTA
  int wc;                   // chunks per activation window
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cta_barrier() { asm volatile("bar.sync 0;" ::: "memory"); }

// A warp's walk over its units: round by round, inside a round a contiguous range of the round's tile-major units.
struct UnitCursor {
  int rd;        // round (window) index; == rounds when the walk is over
  int u, u1;     // position and end of this warp's range inside the round
  int tile, c;   // current unit: row tile (relative to the CTA's first) and chunk
  int cend;      // first chunk after the round's window
};

// The rare steps of the walk are out of line: the unit loop is unrolled over the ring and must stay small (instruction cache).
// Position a cursor on warp `warp`'s range of round rd, or of the first later round in which it has one.
__device__ __noinline__ UnitCursor midm_enter_round(int rd, int rounds, int wc, int chunks, int my_tiles, int warp) {
  UnitCursor q;
  q.u = q.u1 = q.tile = q.c = q.cend = 0;
  for (; rd < rounds; ++rd) {
    const int cw = rd * wc, wcr = min(wc, chunks - cw), ur = my_tiles * wcr;
    const int a = ur * warp / SM_WARPS, b = ur * (warp + 1) / SM_WARPS;
    if (b > a) {
      q.rd = rd; q.u = a; q.u1 = b; q.tile = a / wcr; q.c = cw + a - q.tile * wcr; q.cend = cw + wcr;
      return q;
    }
  }
  q.rd = rounds;
  return q;
}

// window rd -> buffer rd & 1: every thread copies its share (16-byte pieces of the fragment-order activation, chunk sums)
template <typename T>
__device__ __noinline__ void midm_stage(const T* __restrict__ xs_pre, const float* __restrict__ s_pre, uint32_t xs_u32, uint32_t S_u32,
                                        int M, int K, int chunks, int wc, int mpad, int rd, int tid) {
  const int buf = rd & 1, cw = rd * wc, wcr = min(wc, chunks - cw);
  const int pitch = wc * SM_CHUNK + 32;
  const int vec = wcr * 16;
  const T* src0 = xs_pre + (size_t)cw * SM_CHUNK;
  const uint32_t dst0 = xs_u32 + (uint32_t)(buf * mpad * pitch) * (uint32_t)sizeof(T);
  for (int idx = tid; idx < M * vec; idx += SM_THREADS) {
    const int n = idx / vec, v = idx - n * vec;
    cp_async16(dst0 + (uint32_t)(n * pitch + v * 8) * (uint32_t)sizeof(T), src0 + (size_t)n * K + v * 8);
  }
  const uint32_t sdst = S_u32 + (uint32_t)(buf * wc * mpad) * 4u;
  for (int idx = tid; idx < M * wcr; idx += SM_THREADS) {
    const int n = idx / wcr, cc = idx - n * wcr;
    cp_async4(sdst + (uint32_t)(cc * mpad + n) * 4u, s_pre + (size_t)n * chunks + cw + cc);
  }
}


// shared memory: xs[2][MPAD][wc*128 + 32] 16-bit (row pitch + 64 B: the two batch columns of a quarter-warp phase on
// different banks; rows >= M stay zero) | S[2][wc][MPAD] fp32 chunk sums | acc[tiles_per_cta][16][MPAD] fp32 row sums
template <typename T, int NBK>
__global__ void __launch_bounds__(SM_THREADS, 1) w4a16_midm_kernel(const MidMParams p) {
  constexpr int MPAD = 8 * NBK;
  constexpr int CH = 2;                                       // accumulator chains per batch block
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int K = p.K, N = p.N, wc = p.wc;
  const int chunks = K / SM_CHUNK;
  const int rounds = (chunks + wc - 1) / wc;
  const int pitch = wc * SM_CHUNK + 32;                       // elements
  T* xs = reinterpret_cast<T*>(smem_raw);
  float* S = reinterpret_cast<float*>(smem_raw + (size_t)2 * MPAD * pitch * sizeof(T));
  float* acc = S + 2 * wc * MPAD;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // broadcast: the compiler keeps what derives from it uniform
  const int r = lane >> 2, t = lane & 3;

  const int n_tiles = N / 16;
  const int tile0 = min(n_tiles, (int)blockIdx.x * p.tiles_per_cta);
  const int my_tiles = min(n_tiles, tile0 + p.tiles_per_cta) - tile0;
  const int kw = K / 8;                                      // words per row

  auto enter_round = [&](UnitCursor& q, int rd) { q = midm_enter_round(rd, rounds, wc, chunks, my_tiles, warp); };
  auto advance = [&](UnitCursor& q) {
    if (++q.u < q.u1) {
      if (++q.c == q.cend) { q.c = q.rd * wc; ++q.tile; }
    } else {
      enter_round(q, q.rd + 1);
    }
  };
  const uint32_t xs_u32 = smem_addr(xs), S_u32 = smem_addr(S);
  auto stage = [&](int rd) {
    midm_stage<T>(reinterpret_cast<const T*>(p.xs_pre), p.s_pre, xs_u32, S_u32, p.M, K, chunks, wc, MPAD, rd, tid);
  };
  // all of window rd has landed and nobody still reads the other buffer: refill it with window rd + 1
  auto round_entry = [&](int rd) {
    cp_async_wait_all();
    cta_barrier();
    if (rd + 1 < rounds) stage(rd + 1);
  };

  // ---- register ring of MM_PF units (weights of rows r / r+8 and one metadata word per lane: lanes t = 0, 1 carry
  // the scales of rows r, r+8, lanes t = 2, 3 their zero points); the first loads go out before the prologue
  const int gshift = 31 - __clz(p.group / SM_CHUNK);          // chunk -> group
  const size_t row8 = (size_t)8 * kw;
  const bool is_scale = t < 2;
  UnitCursor L, C;
  enter_round(L, 0);
  C = L;
  uint4 qa[MM_PF], qb[MM_PF];
  uint32_t qm[MM_PF];
  auto issue = [&](uint4& a, uint4& b, uint32_t& m) {
    if (L.rd < rounds) {
      const int row = (tile0 + L.tile) * 16 + r;
      const uint32_t* wp = p.qw + (size_t)row * kw + L.c * 16 + t * 4;
      a = ldg_stream(wp);
      b = ldg_stream(wp + row8);
      const size_t mo = (size_t)(L.c >> gshift) * N + row + 8 * (t & 1);
      if (is_scale) m = reinterpret_cast<const unsigned short*>(p.scale_t)[mo];
      else m = p.qzero_t[mo];
      advance(L);
    }
  };
#pragma unroll
  for (int s = 0; s < MM_PF; ++s) { qa[s] = make_uint4(0, 0, 0, 0); qb[s] = qa[s]; qm[s] = 0; issue(qa[s], qb[s], qm[s]); }

  // ---- prologue: zero the batch rows that do not exist (both buffers), their chunk sums and the row sums; window 0
  for (int n = p.M; n < MPAD; ++n)
    for (int idx = tid; idx < 2 * (pitch / 8); idx += SM_THREADS) {
      const int b = idx / (pitch / 8), v = idx - b * (pitch / 8);
      *reinterpret_cast<uint4*>(xs + ((size_t)b * MPAD + n) * pitch + v * 8) = make_uint4(0, 0, 0, 0);
    }
  for (int idx = tid; idx < 2 * wc * MPAD; idx += SM_THREADS)
    if ((idx & (MPAD - 1)) >= p.M) S[idx] = 0.f;
  for (int idx = tid; idx < p.tiles_per_cta * 16 * MPAD; idx += SM_THREADS) acc[idx] = 0.f;
  stage(0);

  float ya[NBK][2], yb[NBK][2];                               // rows r / r+8, batch columns 8j + 2t, 8j + 2t + 1
#pragma unroll
  for (int j = 0; j < NBK; ++j) { ya[j][0] = ya[j][1] = yb[j][0] = yb[j][1] = 0.f; }
  const int lane4 = lane & ~3;
  int cround = -1;                                            // last round this warp has entered

  if (C.rd < rounds) {
    for (;;) {
#pragma unroll
      for (int s = 0; s < MM_PF; ++s) {
        while (cround < C.rd) { ++cround; round_entry(cround); }
        const int buf = C.rd & 1, cl = C.c - C.rd * wc;
        const uint32_t x_addr = smem_addr(xs + ((size_t)buf * MPAD + r) * pitch + cl * SM_CHUNK + t * 8);
        const uint32_t s_addr = smem_addr(S + (size_t)(buf * wc + cl) * MPAD + 2 * t);
        // metadata: the loading lane converts (scale -> fp32, zero point -> OFF + z), then the quad shares it
        float mv;
        {
          const unsigned short hs = (unsigned short)qm[s];
          const float as_scale = Fmt<T>::to_f(*reinterpret_cast<const T*>(&hs));
          const float as_zero = Fmt<T>::OFF + (float)qm[s];
          mv = is_scale ? as_scale : as_zero;
        }
        const float fa = __shfl_sync(0xffffffffu, mv, lane4), fb = __shfl_sync(0xffffffffu, mv, lane4 + 1);
        const float oa = __shfl_sync(0xffffffffu, mv, lane4 + 2), ob = __shfl_sync(0xffffffffu, mv, lane4 + 3);
        // ---- 8 MMAs per batch block over this chunk
        float d[NBK][CH][4];
#pragma unroll
        for (int j = 0; j < NBK; ++j)
#pragma unroll
          for (int h = 0; h < CH; ++h) { d[j][h][0] = d[j][h][1] = d[j][h][2] = d[j][h][3] = 0.f; }
        const uint32_t wra[4] = {qa[s].x, qa[s].y, qa[s].z, qa[s].w}, wrb[4] = {qb[s].x, qb[s].y, qb[s].z, qb[s].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t ua[4], ub[4];
          Fmt<T>::unpack(wra[i], ua);
          Fmt<T>::unpack(wrb[i], ub);
#pragma unroll
          for (int j = 0; j < NBK; ++j) {
            uint4 xb;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(xb.x), "=r"(xb.y), "=r"(xb.z), "=r"(xb.w)
                         : "r"(x_addr + (uint32_t)(j * 8 * pitch) * (uint32_t)sizeof(T) + i * 64));
            Fmt<T>::mma(d[j][0], ua[0], ub[0], ua[1], ub[1], xb.x, xb.y);
            Fmt<T>::mma(d[j][1], ua[2], ub[2], ua[3], ub[3], xb.z, xb.w);
          }
        }
        issue(qa[s], qb[s], qm[s]);                           // refill this stage: unit u + MM_PF
        // ---- chunk epilogue: remove offset and zero point, scale
#pragma unroll
        for (int j = 0; j < NBK; ++j) {
          float2 Sc;
          asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(Sc.x), "=f"(Sc.y) : "r"(s_addr + j * 32));
          float e0 = d[j][0][0], e1 = d[j][0][1], e2 = d[j][0][2], e3 = d[j][0][3];
          e0 += d[j][1][0]; e1 += d[j][1][1]; e2 += d[j][1][2]; e3 += d[j][1][3];
          ya[j][0] = fmaf(fa, fmaf(-oa, Sc.x, e0), ya[j][0]);
          ya[j][1] = fmaf(fa, fmaf(-oa, Sc.y, e1), ya[j][1]);
          yb[j][0] = fmaf(fb, fmaf(-ob, Sc.x, e2), yb[j][0]);
          yb[j][1] = fmaf(fb, fmaf(-ob, Sc.y, e3), yb[j][1]);
        }
        // ---- leaving the row tile (or the walk): add the row sums into the CTA's tile
        const int tile_now = C.tile;
        advance(C);
        const bool over = C.rd == rounds;
        if (over || C.tile != tile_now) {
          float* at = acc + (size_t)tile_now * 16 * MPAD;
#pragma unroll
          for (int j = 0; j < NBK; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int n = 8 * j + 2 * t + e;
              if (n < p.M) {
                atomicAdd(at + r * MPAD + n, ya[j][e]);
                atomicAdd(at + (r + 8) * MPAD + n, yb[j][e]);
              }
              ya[j][e] = 0.f;
              yb[j][e] = 0.f;
            }
        }
        if (over) goto walked;
      }
    }
  }
walked:
  // the rounds this warp had no units in still need its share of the copies and its arrival at the barrier
  while (cround < rounds - 1) { ++cround; round_entry(cround); }
  cta_barrier();
  // ---- convert and store the CTA's tiles: consecutive threads along N
  {
    T* y = reinterpret_cast<T*>(p.y);
    const int rows = my_tiles * 16;
    for (int idx = tid; idx < rows * p.M; idx += SM_THREADS) {
      const int n = idx / rows, rr = idx - n * rows;
      y[(size_t)n * N + (size_t)tile0 * 16 + rr] = Fmt<T>::from_f(acc[(size_t)rr * MPAD + n]);
    }
  }
}


struct MidMPlan { int nbk, wc, grid, tiles_per_cta; size_t smem; };

// batch blocks, window length, grid: whole tiles per CTA, at most one CTA per SM; windows of 16 KB per batch block and buffer
MidMPlan midm_plan(int M, int N, int K) {
  MidMPlan pl;
  pl.nbk = M <= 8 ? 1 : 2;
  const int chunks = K / SM_CHUNK;
  const int wc_max = pl.nbk == 1 ? 32 : 16;
  pl.wc = chunks < wc_max ? chunks : wc_max;
  const int n_tiles = N / 16;
  pl.tiles_per_cta = (n_tiles + sm_count() - 1) / sm_count();
  pl.grid = (n_tiles + pl.tiles_per_cta - 1) / pl.tiles_per_cta;
  const int mpad = 8 * pl.nbk;
  pl.smem = (size_t)2 * mpad * (pl.wc * SM_CHUNK + 32) * 2 + (size_t)2 * pl.wc * mpad * 4 + (size_t)pl.tiles_per_cta * 16 * mpad * 4;
  return pl;
}

template <typename T, int NBK>
int launch_midm(const MidMParams& p, const MidMPlan& pl, cudaStream_t st) {
  static PerDeviceInt smem_set;   // function attributes are per device
  if ((int)pl.smem > smem_set.get()) {
    if (cudaFuncSetAttribute(w4a16_midm_kernel<T, NBK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    smem_set.set((int)pl.smem);
  }
  GQ_LAUNCHED();
  w4a16_midm_kernel<T, NBK><<<pl.grid, SM_THREADS, pl.smem, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

template <typename T>
int run_midm(MidMParams p, const void* x, const int32_t* perm, void* workspace, cudaStream_t st) {
  const MidMPlan pl = midm_plan(p.M, p.N, p.K);
  p.tiles_per_cta = pl.tiles_per_cta;
  p.wc = pl.wc;
  // the gather through perm into fragment order and the chunk sums, once (same workspace layout as w4a16_smallm.cu)
  T* xs_pre = reinterpret_cast<T*>(workspace);
  float* s_pre = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + ((((size_t)p.M * p.K * 2) + 255) & ~(size_t)255));
  p.xs_pre = xs_pre;
  p.s_pre = s_pre;
  dim3 pg((p.K / 8 + 255) / 256, p.M);
  GQ_LAUNCHED();
  w4a16_smallm_xprep_kernel<T><<<pg, 256, 0, st>>>(reinterpret_cast<const T*>(x), perm, xs_pre, s_pre, p.K);
  if (cudaGetLastError() != cudaSuccess) return GQ_ERR_CUDA;
  return pl.nbk == 1 ? launch_midm<T, 1>(p, pl, st) : launch_midm<T, 2>(p, pl, st);
}

}  // namespace

// Does the windowed batch <= 16 kernel take this shape?
bool w4a16_midm_supports(int M, int N, int K, int group) {
  if (M < 1 || M > MM_MAXM || N < 16 || (N & 15) || K < SM_CHUNK || (K % SM_CHUNK) || group < SM_CHUNK ||
      (group & (group - 1)) || (K % group))
    return false;
  return midm_plan(M, N, K).smem <= 200 * 1024;
}

// workspace: M*K*2 B rounded to 256 + M*(K/128)*4 B (inside gq_w4a16_workspace_bytes)
int w4a16_midm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm, void* y,
               void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (!w4a16_midm_supports(M, N, K, group)) return GQ_ERR_SHAPE;
  if (!workspace) return GQ_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(qw) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 255) ||
      (perm && (reinterpret_cast<uintptr_t>(perm) & 15)))
    return GQ_ERR_ALIGN;
  MidMParams p;
  p.qw = qw; p.scale_t = scale_t; p.qzero_t = qzero_t; p.xs_pre = nullptr; p.s_pre = nullptr; p.y = y;
  p.M = M; p.N = N; p.K = K; p.group = group; p.tiles_per_cta = 0; p.wc = 0;
  if (fmt == 0) return run_midm<__half>(p, x, perm, workspace, st);
  if (fmt == 1) return run_midm<__nv_bfloat16>(p, x, perm, workspace, st);
  return GQ_ERR_DTYPE;
}

}  // namespace gq
