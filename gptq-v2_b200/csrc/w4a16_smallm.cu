// K6, batch 1..8:  y (M, N) = x (M, K) . W_hat^T   with W_hat = (q - z) s  from packed int4, one pass over the weights.
//
// The op is a byte stream: 0.5 B per weight from HBM and, at batch <= 8, at most 16 MACs per weight.  What limits a
// CUDA-core GEMV here is instruction issue (unpack + convert + FMA per weight), what limits the tcgen05 path is the
// 128-row MMA floor (w4a16_mma.cu).  This kernel keeps the stream on plain coalesced 16-byte loads and spends
// ~1 issue slot per weight:
//
//   * a packed word becomes four half2 / bf162 registers with one shift + one LOP3 each: the nibble is OR-ed into the
//     mantissa of a constant whose ulp lines up with it (fp16: 64 + q, bf16: 128 + q, both exact);
//   * those registers ARE the A fragment of a warp-level mma.m16n8k16 (16 weight rows x 16 k); the B fragment is the
//     activation, so up to 8 batch columns ride along for free and the products accumulate in fp32;
//   * the constant offset and the zero point come out per 128-column chunk through the chunk's activation sum S:
//         sum_k (q - z) x  =  sum_k (off + q) x  -  (off + z) S,     then  y += s * (...)     (fp32)
//     off is small (64 / 128) so that the cancellation costs < 1e-4 of the chunk's own magnitude.
//
// Work split: units of (16 rows x 128 k).  A CTA (one per SM) owns whole 16-row tiles; its units are dealt to its 16
// warps as equal CONTIGUOUS ranges (stream-K inside the CTA), so a warp carries its fp32 row sums in registers and
// adds them into the tile's shared-memory sums only when its range leaves a tile; the warp that completes a tile
// converts and stores it: no global atomics, no zero-fill, no second pass.  The k order inside a chunk is free (it
// only has to agree between A and B), so lane (r, t) simply takes the 16 contiguous bytes t of rows r and r + 8.
//
// Needs group % 128 == 0, K % 128 == 0, N % 16 == 0, M <= 8; the dispatcher sends everything else elsewhere.
// Reference semantics: reconstruct_nn_linear (quant.py:262-279) followed by nn.Linear on the same input.
#include "gptq_b200_internal.h"
#include "w4a16_frag.cuh"

namespace gq {
namespace {

constexpr int SM_MAXM = 8;
constexpr int SM_PF = 6;                     // units in flight per warp (register ring)

struct SmallMParams {
  const void* x;            // (M, K)
  const uint32_t* qw;       // (N, K/8)
  const void* scale_t;      // (G, N)
  const uint8_t* qzero_t;   // (G, N)
  const int32_t* perm;      // (K) or null
  const void* xs_pre;       // PRE: (M, K) activation already in fragment order (w4a16_smallm_xprep_kernel)
  const float* s_pre;       // PRE: (M, K/128) chunk sums
  void* y;                  // (M, N) 16-bit
  int M, N, K, group;
  int tiles_per_cta;        // row tiles (16 rows) owned by one CTA
};

// shared memory: xs[MP][K + 32] 16-bit (row pitch + 64 B keeps the two batch columns of a quarter-warp phase on
// different banks) | S[K/128][8] fp32 chunk sums | acc[tiles_per_cta][16][MP] fp32 row sums | done[tiles_per_cta] int |
// 256 B of zeros (16 B aligned): the B fragment of the lanes whose batch column does not exist
template <typename T, int MP, bool PRE>
__global__ void __launch_bounds__(SM_THREADS, 1) w4a16_smallm_kernel(const SmallMParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int K = p.K, N = p.N;
  const int chunks = K / SM_CHUNK;
  const int pitch = K + 32;                                  // elements
  T* xs = reinterpret_cast<T*>(smem_raw);
  float* S = reinterpret_cast<float*>(smem_raw + (size_t)MP * pitch * sizeof(T));
  float* acc = S + chunks * 8;
  int* done = reinterpret_cast<int*>(acc + p.tiles_per_cta * 16 * MP);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // broadcast: the compiler keeps what derives from it uniform
  const int r = lane >> 2, t = lane & 3;
  const uint32_t zero_blk = (smem_addr(done + p.tiles_per_cta) + 15u) & ~15u;

  // ---- this CTA owns whole row tiles; its (tile, chunk) units are dealt to the warps as equal contiguous ranges
  const int n_tiles = N / 16;
  const int tile0 = min(n_tiles, (int)blockIdx.x * p.tiles_per_cta);
  const int my_tiles = min(n_tiles, tile0 + p.tiles_per_cta) - tile0;
  const int units = my_tiles * chunks;
  const int u0 = (int)((long long)units * warp / SM_WARPS);
  int left = (int)((long long)units * (warp + 1) / SM_WARPS) - u0;   // units still to compute
  int tile = tile0 + u0 / chunks, c = u0 % chunks;
  const int kw = K / 8;                                      // words per row

  // ---- register ring of SM_PF units (weights of rows r / r+8 and one metadata word per lane: lanes t = 0, 1 carry
  // the scales of rows r, r+8, lanes t = 2, 3 their zero points); the first loads go out before the prologue.
  // Load cursor: running pointers, advanced by one chunk per issue.
  const int gshift = 31 - __clz(p.group / SM_CHUNK);          // chunk -> group
  const int G = K / p.group;
  uint4 qa[SM_PF], qb[SM_PF];
  uint32_t qm[SM_PF];
  int lleft = left, lc = c;
  const uint32_t* wp = p.qw + ((size_t)tile * 16 + r) * kw + c * 16 + t * 4;
  const size_t row8 = (size_t)8 * kw;
  const size_t wrap_w = (size_t)15 * kw + 16;                 // last chunk of a row tile -> first chunk of the next
  const bool is_scale = t < 2;
  // byte cursor into scale_t (2 B) or qzero_t (1 B)
  const unsigned char* mp = is_scale
      ? reinterpret_cast<const unsigned char*>(p.scale_t) + 2 * ((size_t)(c >> gshift) * N + tile * 16 + r + 8 * (t & 1))
      : p.qzero_t + ((size_t)(c >> gshift) * N + tile * 16 + r + 8 * (t & 1));
  const long long mstep = is_scale ? 2ll * N : (long long)N;                      // next group
  const long long mwrap = (is_scale ? 2ll : 1ll) * (16 - (long long)(G - 1) * N);  // last group -> next row tile
  const int gmask = (1 << gshift) - 1;
  auto issue = [&](uint4& a, uint4& b, uint32_t& m) {
    if (lleft > 0) {
      a = ldg_stream(wp);
      b = ldg_stream(wp + row8);
      if (is_scale) m = *reinterpret_cast<const unsigned short*>(mp);
      else m = *mp;
      --lleft;
      if (++lc == chunks) { lc = 0; wp += wrap_w; mp += mwrap; }
      else { wp += 16; if ((lc & gmask) == 0) mp += mstep; }
    }
  };
#pragma unroll
  for (int s = 0; s < SM_PF; ++s) { qa[s] = make_uint4(0, 0, 0, 0); qb[s] = qa[s]; qm[s] = 0; issue(qa[s], qb[s], qm[s]); }

  // ---- prologue: the activation in fragment order + chunk sums into shared memory; clear the row sums
  for (int idx = tid; idx < p.tiles_per_cta * (16 * MP + 1) + 68; idx += SM_THREADS) acc[idx] = 0.f;   // acc, done (0 bits), zero block
  if (PRE) {
    const int vec = K / 8;                                   // uint4 per column
    for (int idx = tid; idx < MP * vec; idx += SM_THREADS) {
      const int n = idx / vec, v = idx - n * vec;
      uint4 val = make_uint4(0, 0, 0, 0);
      if (n < p.M) val = reinterpret_cast<const uint4*>(p.xs_pre)[(size_t)n * vec + v];
      *reinterpret_cast<uint4*>(xs + (size_t)n * pitch + v * 8) = val;
    }
    for (int idx = tid; idx < chunks * 8; idx += SM_THREADS) {
      const int cc = idx >> 3, n = idx & 7;
      S[idx] = n < p.M ? p.s_pre[(size_t)n * chunks + cc] : 0.f;
    }
  } else {
    const T* x = reinterpret_cast<const T*>(p.x);
    const int slots = K / 8;                                 // 8 elements per slot, 16 slots per chunk
    const int total = MP * slots;                            // multiple of 16
    for (int base = 0; base < total; base += 2 * SM_THREADS) {   // warp-uniform trip count (shuffles below), 2 slots in flight
      uint4 out[2];
      float sum[2];
      int nn[2], sl[2];
      bool valid[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int idx = base + h * SM_THREADS + tid;
        valid[h] = idx < total;
        nn[h] = valid[h] ? idx / slots : 0;
        sl[h] = valid[h] ? idx - nn[h] * slots : 0;
        out[h] = make_uint4(0, 0, 0, 0);
        sum[h] = 0.f;
        if (valid[h] && nn[h] < p.M)
          gather_slot<T>(x + (size_t)nn[h] * K, p.perm, (sl[h] >> 4) * SM_CHUNK + ((sl[h] >> 2) & 3) * 32 + (sl[h] & 3) * 8, out[h], sum[h]);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int cc = sl[h] >> 4, tt = (sl[h] >> 2) & 3, ii = sl[h] & 3;
        if (valid[h]) *reinterpret_cast<uint4*>(xs + (size_t)nn[h] * pitch + cc * SM_CHUNK + ii * 32 + tt * 8) = out[h];
        float sm = sum[h];                                   // 16 consecutive idx = one chunk of one column
        sm += __shfl_xor_sync(0xffffffffu, sm, 8);
        sm += __shfl_xor_sync(0xffffffffu, sm, 4);
        sm += __shfl_xor_sync(0xffffffffu, sm, 2);
        sm += __shfl_xor_sync(0xffffffffu, sm, 1);
        if (valid[h] && (sl[h] & 15) == 0) S[cc * 8 + nn[h]] = sm;
      }
    }
    if (MP < 8)
      for (int idx = tid; idx < chunks * 8; idx += SM_THREADS)
        if ((idx & 7) >= MP) S[idx] = 0.f;
  }
  __syncthreads();
  if (left <= 0) return;

  // B fragment source: lane (n = r, t) carries batch column r; lanes without a column read the zero block (stride 0)
  const bool has_x = r < MP;
  const uint32_t x_base = has_x ? smem_addr(xs + (size_t)r * pitch + t * 8) : zero_blk;
  const uint32_t x_step = has_x ? SM_CHUNK * (uint32_t)sizeof(T) : 0u;
  uint32_t x_addr = x_base + (uint32_t)c * x_step;
  const uint32_t s_base = smem_addr(S + 2 * t);
  uint32_t s_addr = s_base + (uint32_t)c * 32u;
  float ya[2] = {0.f, 0.f}, yb[2] = {0.f, 0.f};               // rows r / r+8, batch columns 2t, 2t+1
  const int lane4 = lane & ~3;
  int run = 0;                                                // chunks accumulated since the last flush

  for (;;) {
#pragma unroll
    for (int s = 0; s < SM_PF; ++s) {
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
      // ---- 8 MMAs over this chunk, two independent accumulator chains
      float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
      const uint32_t wra[4] = {qa[s].x, qa[s].y, qa[s].z, qa[s].w}, wrb[4] = {qb[s].x, qb[s].y, qb[s].z, qb[s].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 xb;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(xb.x), "=r"(xb.y), "=r"(xb.z), "=r"(xb.w) : "r"(x_addr + i * 64));
        uint32_t ua[4], ub[4];
        Fmt<T>::unpack(wra[i], ua);
        Fmt<T>::unpack(wrb[i], ub);
        Fmt<T>::mma(d0, ua[0], ub[0], ua[1], ub[1], xb.x, xb.y);
        Fmt<T>::mma(d1, ua[2], ub[2], ua[3], ub[3], xb.z, xb.w);
      }
      issue(qa[s], qb[s], qm[s]);                             // refill this stage: unit u + SM_PF
      // ---- chunk epilogue: remove offset and zero point, scale
      {
        float2 Sc;
        asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(Sc.x), "=f"(Sc.y) : "r"(s_addr));
        ya[0] = fmaf(fa, fmaf(-oa, Sc.x, d0[0] + d1[0]), ya[0]);
        ya[1] = fmaf(fa, fmaf(-oa, Sc.y, d0[1] + d1[1]), ya[1]);
        yb[0] = fmaf(fb, fmaf(-ob, Sc.x, d0[2] + d1[2]), yb[0]);
        yb[1] = fmaf(fb, fmaf(-ob, Sc.y, d0[3] + d1[3]), yb[1]);
      }
      ++run;
      // ---- leave the row tile (or the range): add the row sums into the CTA's tile; the warp that completes a
      // tile converts and stores it
      const bool last = --left == 0;
      const bool wrap = c + 1 == chunks;
      if (last || wrap) {
        const int tl = tile - tile0;
        float* at = acc + (size_t)tl * 16 * MP;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int n = 2 * t + e;
          if (n < MP) {
            atomicAdd(at + r * MP + n, ya[e]);
            atomicAdd(at + (r + 8) * MP + n, yb[e]);
          }
        }
        ya[0] = ya[1] = yb[0] = yb[1] = 0.f;
        __syncwarp();
        int fin = 0;
        if (lane == 0) {
          __threadfence_block();
          fin = atomicAdd(done + tl, run) + run == chunks;
        }
        run = 0;
        fin = __shfl_sync(0xffffffffu, fin, 0);
        if (fin) {
          __threadfence_block();
          T* y = reinterpret_cast<T*>(p.y);
          for (int idx = lane; idx < 16 * p.M; idx += 32) {
            const int row = idx & 15, n = idx >> 4;
            y[(size_t)n * N + tile * 16 + row] = Fmt<T>::from_f(*reinterpret_cast<volatile float*>(at + row * MP + n));
          }
        }
        if (last) return;
      }
      if (wrap) { c = 0; ++tile; x_addr = x_base; s_addr = s_base; } else { ++c; x_addr += x_step; s_addr += 32u; }
    }
  }
}

template <typename T, int MP, bool PRE>
int launch_smallm(const SmallMParams& p, size_t smem, int grid, cudaStream_t st) {
  static PerDeviceInt smem_set;   // function attributes are per device
  if ((int)smem > smem_set.get()) {
    if (cudaFuncSetAttribute(w4a16_smallm_kernel<T, MP, PRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    smem_set.set((int)smem);
  }
  GQ_LAUNCHED();
  w4a16_smallm_kernel<T, MP, PRE><<<grid, SM_THREADS, smem, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

inline int pad_m(int M) { return M <= 1 ? 1 : M <= 2 ? 2 : M <= 4 ? 4 : 8; }


// grid and row tiles per CTA: whole tiles per CTA, at most one CTA per SM
void split_tiles(int N, int& grid, int& tiles_per_cta) {
  const int n_tiles = N / 16;
  tiles_per_cta = (n_tiles + sm_count() - 1) / sm_count();
  grid = (n_tiles + tiles_per_cta - 1) / tiles_per_cta;
}

size_t smallm_smem(int M, int N, int K) {
  int grid, tpc;
  split_tiles(N, grid, tpc);
  const int mp = pad_m(M);
  return (size_t)mp * (K + 32) * 2 + (size_t)(K / SM_CHUNK) * 32 + (size_t)tpc * (16 * mp + 1) * 4 + 272;
}

constexpr int SM_PRE_ELEMS = 16384;          // M_pad * K from which the gather is hoisted into its own launch instead of being repeated
                                             // by every CTA (measured: batch 8 on 4096 x 4096 12.9 -> 8.5 us, batch 4 9.4 -> 8.2 us, batch 2 on
                                             // K = 11008 14.7 -> 12.6 us; from 8192 down the extra launch costs more than it saves)

template <typename T>
int run_smallm(SmallMParams p, void* workspace, cudaStream_t st) {
  const int mp = pad_m(p.M);
  int grid;
  split_tiles(p.N, grid, p.tiles_per_cta);
  const size_t smem = smallm_smem(p.M, p.N, p.K);
  const bool pre = (long long)mp * p.K >= SM_PRE_ELEMS;
  if (pre) {
    T* xs_pre = reinterpret_cast<T*>(workspace);
    float* s_pre = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + ((((size_t)p.M * p.K * 2) + 255) & ~(size_t)255));
    p.xs_pre = xs_pre;
    p.s_pre = s_pre;
    dim3 pg((p.K / 8 + 255) / 256, p.M);
    GQ_LAUNCHED();
    w4a16_smallm_xprep_kernel<T><<<pg, 256, 0, st>>>(reinterpret_cast<const T*>(p.x), p.perm, xs_pre, s_pre, p.K);
  }
#define GQ_SMALLM_CASE(MPV)                                                               \
  case MPV: return pre ? launch_smallm<T, MPV, true>(p, smem, grid, st) : launch_smallm<T, MPV, false>(p, smem, grid, st);
  switch (mp) {
    GQ_SMALLM_CASE(1)
    GQ_SMALLM_CASE(2)
    GQ_SMALLM_CASE(4)
    default: return pre ? launch_smallm<T, 8, true>(p, smem, grid, st) : launch_smallm<T, 8, false>(p, smem, grid, st);
  }
#undef GQ_SMALLM_CASE
}

}  // namespace

// Does the batch-1..8 kernel take this shape?  (shared memory: the whole activation, M_pad x (K + 32) x 2 B)
bool w4a16_smallm_supports(int M, int N, int K, int group) {
  if (M < 1 || M > SM_MAXM || N < 16 || (N & 15) || K < SM_CHUNK || (K % SM_CHUNK) || group < SM_CHUNK ||
      (group & (group - 1)) || (K % group))
    return false;
  return smallm_smem(M, N, K) <= 200 * 1024;
}

// workspace (only used when M_pad * K >= 16384): M*K*2 B rounded to 256 + M*(K/128)*4 B
int w4a16_smallm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm, void* y,
                 void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st) {
  if (!w4a16_smallm_supports(M, N, K, group)) return GQ_ERR_SHAPE;
  if (!workspace) return GQ_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(qw) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 255) ||
      (perm && (reinterpret_cast<uintptr_t>(perm) & 15)))
    return GQ_ERR_ALIGN;
  SmallMParams p;
  p.x = x; p.qw = qw; p.scale_t = scale_t; p.qzero_t = qzero_t; p.perm = perm;
  p.xs_pre = nullptr; p.s_pre = nullptr; p.y = y;
  p.M = M; p.N = N; p.K = K; p.group = group; p.tiles_per_cta = 0;
  if (fmt == 0) return run_smallm<__half>(p, workspace, st);
  if (fmt == 1) return run_smallm<__nv_bfloat16>(p, workspace, st);
  return GQ_ERR_DTYPE;
}

}  // namespace gq
