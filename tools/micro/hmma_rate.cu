// Issue rate of the legacy warp-level mma.sync.m16n8k16 (f16 x f16 -> f32) on one SM, for 1..32 resident warps with
// four independent accumulator chains each.  Prints cycles per HMMA per SM sub-partition.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hmma_rate hmma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void hmma_loop(float* out, long long* cyc, int iters) {
  float d[4][4] = {};
  uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 0x3c003c00u, a3 = 0x3c003c00u, b0 = 0x3c003c00u, b1 = 0x38003800u;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(d[j][0]), "+f"(d[j][1]), "+f"(d[j][2]), "+f"(d[j][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  __syncthreads();
  const long long t1 = clock64();
  float s = 0.f;
  for (int j = 0; j < 4; ++j) for (int k = 0; k < 4; ++k) s += d[j][k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, 1024 * 148 * 4); cudaMalloc(&cyc, 148 * 8);
  const int iters = 4096;
  for (int warps : {1, 2, 4, 8, 16, 32}) {
    hmma_loop<<<148, warps * 32>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    hmma_loop<<<148, warps * 32>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double per_smsp = (double)c / ((double)iters * 4 * warps / (warps < 4 ? warps : 4));
    printf("warps/SM %2d: %lld cycles, %.2f cycles per HMMA.16816 per sub-partition, %.0f dense FMA/clk/SM\n", warps, c, per_smsp,
           16.0 * 8 * 16 * iters * 4 * warps / (double)c);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
