// What read pattern does a one-pass stream over a packed weight matrix (N rows x K/2 bytes) reach on B200?
//   A  the K6 small-batch kernel's pattern (w4a16_smallm.cu): lane (r, t) of a warp loads 16 B of rows r and r + 8, i.e. a
//      warp instruction covers 8 rows x 64 B; a register ring of 6 units (16 rows x 64 B) per warp; 16 warps per CTA walk
//      contiguous unit ranges of the CTA's row tiles
//   B  the same bytes fetched by cp.async.bulk (one elected thread, 16 rows x RUN contiguous bytes per stage into shared
//      memory, mbarrier ring), consumed by all warps with ld.shared.v4
//   C  a plain linear read: every warp instruction covers 512 contiguous bytes
// No arithmetic beyond an XOR (so that the loads are not dead).  Buffers are rotated so that every launch reads from DRAM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_patterns stream_patterns.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint4 ldg_stream(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- A
template <int PF>
__global__ void __launch_bounds__(512, 1) pattern_a(const uint32_t* __restrict__ qw, uint32_t* out, int N, int K, int tiles_per_cta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, r = lane >> 2, t = lane & 3;
  const int chunks = K / 128, kw = K / 8, n_tiles = N / 16;
  const int tile0 = min(n_tiles, (int)blockIdx.x * tiles_per_cta);
  const int my_tiles = min(n_tiles, tile0 + tiles_per_cta) - tile0;
  const int units = my_tiles * chunks;
  const int u0 = (int)((long long)units * warp / 16);
  int left = (int)((long long)units * (warp + 1) / 16) - u0;
  int tile = tile0 + u0 / chunks, c = u0 % chunks;
  const uint32_t* wp = qw + ((size_t)tile * 16 + r) * kw + c * 16 + t * 4;
  const size_t row8 = (size_t)8 * kw, wrap_w = (size_t)15 * kw + 16;
  uint4 qa[PF], qb[PF];
  int lleft = left, lc = c;
  auto issue = [&](uint4& a, uint4& b) {
    if (lleft > 0) {
      a = ldg_stream(wp); b = ldg_stream(wp + row8);
      --lleft;
      if (++lc == chunks) { lc = 0; wp += wrap_w; } else wp += 16;
    }
  };
#pragma unroll
  for (int s = 0; s < PF; ++s) { qa[s] = make_uint4(0, 0, 0, 0); qb[s] = qa[s]; issue(qa[s], qb[s]); }
  uint32_t x = 0;
  while (left > 0) {
#pragma unroll
    for (int s = 0; s < PF; ++s) {
      if (left > 0) {
        x ^= qa[s].x ^ qa[s].y ^ qa[s].z ^ qa[s].w ^ qb[s].x ^ qb[s].y ^ qb[s].z ^ qb[s].w;
        issue(qa[s], qb[s]);
        --left;
      }
    }
  }
  if (x == 0x12345678u) out[threadIdx.x] = x;
}

// ---- B
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// stage = 16 rows x RUN bytes; a CTA owns whole row tiles and walks each along K
template <int RUN, int STAGES>
__global__ void __launch_bounds__(512 + 32, 1) pattern_b(const uint32_t* __restrict__ qw, uint32_t* out, int N, int K, int tiles_per_cta) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bars[2 * STAGES];
  const int n_tiles = N / 16;
  const int tile0 = min(n_tiles, (int)blockIdx.x * tiles_per_cta);
  const int my_tiles = min(n_tiles, tile0 + tiles_per_cta) - tile0;
  const int row_bytes = K / 2;
  const int runs = row_bytes / RUN;                          // K/2 % RUN == 0 assumed
  const int total = my_tiles * runs;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(smem_u32(&bars[s]), 1); mbar_init(smem_u32(&bars[STAGES + s]), 16); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 16) {
    if (lane == 0) {
      for (int i = 0; i < total; ++i) {
        const int s = i % STAGES, ph = (i / STAGES) & 1;
        if (i >= STAGES) mbar_wait(smem_u32(&bars[STAGES + s]), ph ^ 1);
        const int tl = i / runs, ru = i - tl * runs;
        const uint32_t full = smem_u32(&bars[s]);
        mbar_expect_tx(full, 16 * RUN);
        const unsigned char* src = reinterpret_cast<const unsigned char*>(qw) + ((size_t)(tile0 + tl) * 16) * row_bytes + (size_t)ru * RUN;
        for (int rr = 0; rr < 16; ++rr)
          bulk_load_1d(smem_u32(smem + (size_t)s * 16 * RUN + rr * RUN), src + (size_t)rr * row_bytes, RUN, full);
      }
    }
    return;
  }
  uint32_t x = 0;
  for (int i = 0; i < total; ++i) {
    const int s = i % STAGES, ph = (i / STAGES) & 1;
    mbar_wait(smem_u32(&bars[s]), ph);
    const uint4* st = reinterpret_cast<const uint4*>(smem + (size_t)s * 16 * RUN);
    for (int e = threadIdx.x; e < 16 * RUN / 16; e += 512) { const uint4 v = st[e]; x ^= v.x ^ v.y ^ v.z ^ v.w; }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&bars[STAGES + s]));
  }
  if (x == 0x12345678u) out[threadIdx.x] = x;
}

// ---- C
__global__ void __launch_bounds__(512, 1) pattern_c(const uint4* __restrict__ qw, uint32_t* out, size_t n16) {
  const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
  const size_t b = blockIdx.x * per, e = min(n16, b + per);
  uint32_t x = 0;
  for (size_t i = b + threadIdx.x; i < e; i += 512 * 4) {
    uint4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = i + j * 512 < e ? ldg_stream(reinterpret_cast<const uint32_t*>(qw + i + j * 512)) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) x ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
  }
  if (x == 0x12345678u) out[threadIdx.x] = x;
}

template <typename F>
float time_us(F launch, int nbuf, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < nbuf; ++i) launch(i);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int rp = 0; rp < reps; ++rp)
    for (int i = 0; i < nbuf; ++i) launch(i);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  CK(cudaGetLastError());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return ms * 1000.f / (reps * nbuf);
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const int shapes[][2] = {{28672, 8192}, {8192, 28672}, {8192, 8192}, {4096, 4096}};
  uint32_t* out;
  CK(cudaMalloc(&out, 4096));
  for (auto& sh : shapes) {
    const int N = sh[0], K = sh[1];
    const size_t bytes = (size_t)N * K / 2;
    const int nbuf = (int)(700e6 / bytes) + 2;
    std::vector<uint32_t*> bufs(nbuf);
    for (auto& b : bufs) { CK(cudaMalloc(&b, bytes)); CK(cudaMemset(b, 0x5a, bytes)); }
    const int n_tiles = N / 16;
    const int tpc = (n_tiles + sms - 1) / sms, grid = (n_tiles + tpc - 1) / tpc;
    const double gb = bytes / 1e3;   // bytes/us -> GB/s = bytes / us / 1e3
    float ta6 = time_us([&](int i) { pattern_a<6><<<grid, 512>>>(bufs[i], out, N, K, tpc); }, nbuf, 5);
    float ta10 = time_us([&](int i) { pattern_a<10><<<grid, 512>>>(bufs[i], out, N, K, tpc); }, nbuf, 5);
    printf("%5d x %5d  A ring 6: %6.1f us %6.0f GB/s   A ring 10: %6.1f us %6.0f GB/s  (grid %d)\n", N, K, ta6, gb / ta6, ta10, gb / ta10, grid);
    {
      constexpr int RUN = 1024, ST = 8;
      CK(cudaFuncSetAttribute(pattern_b<RUN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * 16 * RUN));
      float tb = time_us([&](int i) { pattern_b<RUN, ST><<<grid, 544, ST * 16 * RUN>>>(bufs[i], out, N, K, tpc); }, nbuf, 5);
      printf("               B bulk 16 x 1024 B x 8 stages: %6.1f us %6.0f GB/s\n", tb, gb / tb);
    }
    {
      constexpr int RUN = 512, ST = 12;
      CK(cudaFuncSetAttribute(pattern_b<RUN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * 16 * RUN));
      float tb = time_us([&](int i) { pattern_b<RUN, ST><<<grid, 544, ST * 16 * RUN>>>(bufs[i], out, N, K, tpc); }, nbuf, 5);
      printf("               B bulk 16 x  512 B x 12 stages: %6.1f us %6.0f GB/s\n", tb, gb / tb);
    }
    {
      constexpr int RUN = 256, ST = 16;
      CK(cudaFuncSetAttribute(pattern_b<RUN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * 16 * RUN));
      float tb = time_us([&](int i) { pattern_b<RUN, ST><<<grid, 544, ST * 16 * RUN>>>(bufs[i], out, N, K, tpc); }, nbuf, 5);
      printf("               B bulk 16 x  256 B x 16 stages: %6.1f us %6.0f GB/s\n", tb, gb / tb);
    }
    float tc = time_us([&](int i) { pattern_c<<<sms, 512>>>(reinterpret_cast<const uint4*>(bufs[i]), out, bytes / 16); }, nbuf, 5);
    float tc2 = time_us([&](int i) { pattern_c<<<2 * sms, 512>>>(reinterpret_cast<const uint4*>(bufs[i]), out, bytes / 16); }, nbuf, 5);
    printf("               C linear: %6.1f us %6.0f GB/s  (2 CTAs/SM: %6.1f us %6.0f GB/s)\n", tc, gb / tc, tc2, gb / tc2);
    for (auto& b : bufs) CK(cudaFree(b));
  }
  printf("no error\n");
  return 0;
}
