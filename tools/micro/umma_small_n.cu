// Micro-benchmark: cost of tcgen05.mma M=128, K=16 for small N, A from TMEM (TS) or smem (SS),
// accumulating into ONE buffer (dependent chain) or rotating over R independent buffers.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../gptq-v2_b200/csrc umma_small_n.cu -o umma_small_n
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace gq;

__global__ void __launch_bounds__(128, 1) k(int N, int rot, int ts, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 64 * 1024;
  const uint32_t slot = bar + 16;
  volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (slot - smem_u32(smem_raw)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 16 * 1024; i += 128) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc<1>(slot, 512); tmem_relinquish<1>(); }
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = *slot_ptr;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = umma_idesc(0u, 0u, 0u, 128, (uint32_t)N);
    const uint64_t ad = umma_smem_desc(base, 16, 1024, 2u);
    const uint64_t bd = umma_smem_desc(base + 32 * 1024, 16, 1024, 2u);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tm + (i % rot) * (N < 32 ? 32 : N);
      if (ts) mma_f16_ts(d, tm + 256 + (i & 7) * 8, bd, idesc, 1u);
      else mma_f16<1>(d, ad, bd, idesc, 1u);
    }
    tc_commit_cg1(bar);
    mbar_wait(bar, 0);
    long long t1 = clock64();
    out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  const int iters = 4096;
  for (int ts = 0; ts < 2; ++ts)
    for (int N : {16, 32, 64, 128, 256})
      for (int rot : {1, 2, 4, 8}) {
        if (rot * (N < 32 ? 32 : N) > 256) continue;
        long long h = 0;
        k<<<1, 128, 80 * 1024>>>(N, rot, ts, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("%s N=%3d rot=%d : %7.1f cycles/MMA  (%s)\n", ts ? "TS" : "SS", N, rot, (double)h / iters, cudaGetErrorString(e));
      }
  return 0;
}
