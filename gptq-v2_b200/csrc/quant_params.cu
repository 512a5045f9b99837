// Quantizer.find_params and Quantizer.mse on the solver's transposed (C, R) layout  (SURVEY.md section 8f rank 4).
//
// Reference: quant.py:89-118 (min/max and symmetric grids) and quant.py:152-156 (the shrink-factor grid search,
// use_kernel=False).  One CTA owns 32 weight rows (lane = row, so every global access is a coalesced 128-byte row of
// Wt) of one group; its 8 warps split the group's columns (find_params) or the grid points (mse).
//
//   find_params:  scale = (relu(max) - (-relu(-min))) / maxq + 1e-12, zero = round(-min / scale)          (asymmetric)
//                 scale = absmax * (2 / maxq) + 1e-12,               zero = round((maxq + 1) / 2)          (symmetric)
//   mse_refine:   for every grid point p_i: s_i = scale * p_i, err_i = sum_c |deq(q(x_c; s_i)) - x_c|^norm; scale *= p_argmin
//                 (first minimum wins, like torch.argmin).  The group's 32 x gs tile is staged in shared memory once and
//                 read n_grid times; x / s_i is an IEEE division and rounding is half-to-even, as in quant.py:19.
//
// The EXL2 step that follows in the reference (quant.py:109-112, 157-165) works on the (R, G) scales only; the host
// mirror does it with a handful of torch calls on the device (gptq_py.py of this package).
#include <algorithm>
#include "gptq_b200_internal.h"

namespace gq {
namespace {

constexpr int QP_ROWS = 32;
constexpr int QP_WARPS = 8;

__global__ void __launch_bounds__(QP_ROWS * QP_WARPS) find_params_kernel(const float* __restrict__ Wt, int R, int group,
                                                                          const float* __restrict__ maxq_g, int sym,
                                                                          float* __restrict__ scale_gr,
                                                                          float* __restrict__ qzero_gr) {
  __shared__ float red[2][QP_WARPS][QP_ROWS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * QP_ROWS + lane;
  const int g = blockIdx.y;
  const bool active = r < R;
  float mx = -INFINITY, mn = INFINITY;
  if (active) {
    const float* src = Wt + (size_t)g * group * R + r;
#pragma unroll 8
    for (int c = warp; c < group; c += QP_WARPS) {
      const float v = src[(size_t)c * R];
      mx = fmaxf(mx, v);
      mn = fminf(mn, v);
    }
  }
  red[0][warp][lane] = mx;
  red[1][warp][lane] = mn;
  __syncthreads();
  if (warp != 0 || !active) return;
#pragma unroll
  for (int i = 1; i < QP_WARPS; ++i) {
    mx = fmaxf(mx, red[0][i][lane]);
    mn = fminf(mn, red[1][i][lane]);
  }
  const float maxq = maxq_g[g];
  float s, z;
  if (sym) {
    const float amax = fmaxf(fabsf(mx), fabsf(mn));
    s = __fadd_rn(__fmul_rn(amax, __fdiv_rn(2.f, maxq)), 1e-12f);             // quant.py:105
    z = rintf(__fmul_rn(__fadd_rn(maxq, 1.f), 0.5f));                         // quant.py:106
  } else {
    const float xmax = fmaxf(mx, 0.f);
    const float xmin = -fmaxf(-mn, 0.f);
    s = __fadd_rn(__fdiv_rn(__fsub_rn(xmax, xmin), maxq), 1e-12f);            // quant.py:116
    z = rintf(__fdiv_rn(-xmin, s));                                           // quant.py:117
  }
  scale_gr[(size_t)g * R + r] = s;
  qzero_gr[(size_t)g * R + r] = z;
}

// Correctly rounded a / b from rb = RN(1 / b) (Markstein; the same three operations as the solver's div_by, checked
// against IEEE division in the CPU suite): the divisor is fixed for a whole grid point.
__device__ __forceinline__ float div_by_scale(float a, float b, float rb) {
  const float q = __fmul_rn(a, rb);
  return __fmaf_rn(__fmaf_rn(-q, b, a), rb, q);
}

// err of one grid point for this lane's row over the staged tile
template <bool SQUARE>
__device__ __forceinline__ float grid_error(const float* tile, int group, int lane, float s, float z, float maxq, float norm) {
  float acc = 0.f;
  const float rs = __frcp_rn(s);
#pragma unroll 4
  for (int c = 0; c < group; ++c) {
    const float x = tile[c * QP_ROWS + lane];
    float q = __fadd_rn(rintf(div_by_scale(x, s, rs)), z);
    q = fminf(fmaxf(q, 0.f), maxq);
    const float e = fabsf(__fsub_rn(__fmul_rn(__fsub_rn(q, z), s), x));
    acc = __fadd_rn(acc, SQUARE ? __fmul_rn(e, e) : powf(e, norm));
  }
  return acc;
}

__global__ void __launch_bounds__(QP_ROWS * QP_WARPS) mse_refine_kernel(const float* __restrict__ Wt, int R, int group,
                                                                         const float* __restrict__ maxq_g,
                                                                         const float* __restrict__ grid, int n_grid, float norm,
                                                                         float* __restrict__ scale_gr,
                                                                         const float* __restrict__ qzero_gr,
                                                                         int* __restrict__ index_gr) {
  extern __shared__ __align__(16) float qp_smem[];
  float* tile = qp_smem;                                   // [group][32]
  float* best_e = tile + (size_t)group * QP_ROWS;          // [8][32]
  int* best_i = reinterpret_cast<int*>(best_e + QP_WARPS * QP_ROWS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * QP_ROWS + lane;
  const int g = blockIdx.y;
  const bool active = r < R;
  {
    const float* src = Wt + (size_t)g * group * R + r;
    for (int c = warp; c < group; c += QP_WARPS) tile[c * QP_ROWS + lane] = active ? src[(size_t)c * R] : 0.f;
  }
  __syncthreads();
  const float maxq = maxq_g[g];
  const float s0 = active ? scale_gr[(size_t)g * R + r] : 1.f;
  const float z = active ? qzero_gr[(size_t)g * R + r] : 0.f;
  float be = INFINITY;
  int bi = 0x7fffffff;
  for (int i = warp; i < n_grid; i += QP_WARPS) {          // ascending i: a strict < keeps the first minimum
    const float s = __fmul_rn(s0, grid[i]);
    const float e = norm == 2.f ? grid_error<true>(tile, group, lane, s, z, maxq, norm)
                                : grid_error<false>(tile, group, lane, s, z, maxq, norm);
    if (e < be) { be = e; bi = i; }
  }
  best_e[warp * QP_ROWS + lane] = be;
  best_i[warp * QP_ROWS + lane] = bi;
  __syncthreads();
  if (warp != 0 || !active) return;
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) {
    const float e = best_e[w * QP_ROWS + lane];
    const int i = best_i[w * QP_ROWS + lane];
    if (e < be || (e == be && i < bi)) { be = e; bi = i; }
  }
  if (bi == 0x7fffffff) bi = 0;                            // all-NaN row: torch.argmin returns the first NaN; keep p[0]
  scale_gr[(size_t)g * R + r] = __fmul_rn(s0, grid[bi]);   // quant.py:156
  if (index_gr) index_gr[(size_t)g * R + r] = bi;
}

// Round to nearest with given parameters (gptq_py.py:101-124: Quantizer.quantize / dequantize of every group, nothing
// propagates): q = clamp(round(w / s) + z, 0, maxq), w_hat = (q - z) * s, element-wise on the (C, R) layout.
__global__ void __launch_bounds__(256) rtn_quantize_kernel(const float* __restrict__ Wt, int R, int group, int n_cols,
                                                           const float* __restrict__ maxq_g, const float* __restrict__ scale_gr,
                                                           const float* __restrict__ qzero_gr, float* __restrict__ quant_t,
                                                           uint8_t* __restrict__ codes_t) {
  const int r4 = R >> 2;
  const size_t n = (size_t)n_cols * r4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / r4), r = (int)(i - (size_t)c * r4) * 4;
    const int g = c / group;
    const float maxq = maxq_g[g];
    const float4 w = *reinterpret_cast<const float4*>(Wt + (size_t)c * R + r);
    const float4 s = *reinterpret_cast<const float4*>(scale_gr + (size_t)g * R + r);
    const float4 z = *reinterpret_cast<const float4*>(qzero_gr + (size_t)g * R + r);
    const float wv[4] = {w.x, w.y, w.z, w.w}, sv[4] = {s.x, s.y, s.z, s.w}, zv[4] = {z.x, z.y, z.z, z.w};
    float dq[4];
    uint32_t codes = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float q = __fadd_rn(rintf(__fdiv_rn(wv[k], sv[k])), zv[k]);          // quant.py:19
      q = fminf(fmaxf(q, 0.f), maxq);
      dq[k] = __fmul_rn(__fsub_rn(q, zv[k]), sv[k]);                       // quant.py:30
      codes |= (uint32_t)(int)q << (8 * k);
    }
    if (quant_t) *reinterpret_cast<float4*>(quant_t + (size_t)c * R + r) = make_float4(dq[0], dq[1], dq[2], dq[3]);
    if (codes_t) *reinterpret_cast<uint32_t*>(codes_t + (size_t)c * R + r) = codes;
  }
}

}  // namespace

// n_groups groups of `group` columns each, starting where Wt / quant_t / codes_t / scale_gr / qzero_gr / maxq_groups_dev point.
int rtn_quantize(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* scale_gr,
                 const float* qzero_gr, float* quant_t, uint8_t* codes_t, cudaStream_t st) {
  if (R <= 0 || (R & 3) || group <= 0 || n_groups <= 0) return GQ_ERR_SHAPE;
  if (!Wt || !maxq_groups_dev || !scale_gr || !qzero_gr) return GQ_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(Wt) | reinterpret_cast<uintptr_t>(quant_t) | reinterpret_cast<uintptr_t>(scale_gr) |
       reinterpret_cast<uintptr_t>(qzero_gr)) & 15) return GQ_ERR_ALIGN;
  if (reinterpret_cast<uintptr_t>(codes_t) & 3) return GQ_ERR_ALIGN;
  const size_t n = (size_t)group * n_groups * (R >> 2);
  const int grid = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
  GQ_LAUNCHED();
  rtn_quantize_kernel<<<grid, 256, 0, st>>>(Wt, R, group, group * n_groups, maxq_groups_dev, scale_gr, qzero_gr, quant_t, codes_t);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

// Wt points at the first row of the first group; scale_gr / qzero_gr / maxq_groups_dev at that group's entries.
int find_params(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, int sym, float* scale_gr,
                float* qzero_gr, cudaStream_t st) {
  if (R <= 0 || group <= 0 || n_groups <= 0 || n_groups > 65535) return GQ_ERR_SHAPE;
  if (!Wt || !maxq_groups_dev || !scale_gr || !qzero_gr) return GQ_ERR_ARG;
  dim3 grid((R + QP_ROWS - 1) / QP_ROWS, n_groups);
  GQ_LAUNCHED();
  find_params_kernel<<<grid, QP_ROWS * QP_WARPS, 0, st>>>(Wt, R, group, maxq_groups_dev, sym, scale_gr, qzero_gr);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

int mse_refine(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* grid_dev,
               int n_grid, float norm, float* scale_gr, const float* qzero_gr, int* index_gr, cudaStream_t st) {
  if (R <= 0 || group <= 0 || group > 1024 || n_groups <= 0 || n_groups > 65535 || n_grid <= 0) return GQ_ERR_SHAPE;
  if (!Wt || !maxq_groups_dev || !grid_dev || !scale_gr || !qzero_gr) return GQ_ERR_ARG;
  const int smem = (group * QP_ROWS + 2 * QP_WARPS * QP_ROWS) * 4;
  static PerDeviceInt smem_set;          // function attributes are per device
  if (smem > 48 * 1024 && smem > smem_set.get()) {
    if (cudaFuncSetAttribute(mse_refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return GQ_ERR_CUDA;
    smem_set.set(smem);
  }
  dim3 grid((R + QP_ROWS - 1) / QP_ROWS, n_groups);
  GQ_LAUNCHED();
  mse_refine_kernel<<<grid, QP_ROWS * QP_WARPS, smem, st>>>(Wt, R, group, maxq_groups_dev, grid_dev, n_grid, norm, scale_gr,
                                                            qzero_gr, index_gr);
  return cudaGetLastError() == cudaSuccess ? GQ_OK : GQ_ERR_CUDA;
}

}  // namespace gq
