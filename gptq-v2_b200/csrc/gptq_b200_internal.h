// Internal (C++) declarations shared by the .cu files; the public C ABI is include/gptq_b200.h
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define GQ_OK 0
#define GQ_ERR_CUDA 1        // a CUDA runtime call / launch failed (see cudaGetLastError)
#define GQ_ERR_SHAPE 2       // unsupported shape (see the precondition in the header)
#define GQ_ERR_DTYPE 3
#define GQ_ERR_ALIGN 4       // pointer not 16-byte aligned
#define GQ_ERR_DRIVER 5      // cuTensorMapEncodeTiled entry point not found
#define GQ_ERR_TENSORMAP 6   // cuTensorMapEncodeTiled rejected the descriptor
#define GQ_ERR_NOT_SPD 7     // Cholesky met a non-positive pivot (caller adds damping and retries)
#define GQ_ERR_ARG 8
#define GQ_ERR_LIMIT 9       // size beyond the fixed scratch of an allocation-free entry point

#include <atomic>

namespace gq {

// cudaFuncSetAttribute is per DEVICE: "done once" flags are indexed by the current device (a process may drive several
// GPUs, e.g. an HF device_map) and atomic (llama_driver.py runs worker threads).
constexpr int GQ_MAX_DEVICES = 64;
struct PerDeviceInt {
  std::atomic<int> v[GQ_MAX_DEVICES];
  PerDeviceInt() { for (auto& x : v) x.store(0, std::memory_order_relaxed); }
  static int dev() { int d = 0; cudaGetDevice(&d); return (d >= 0 && d < GQ_MAX_DEVICES) ? d : 0; }
  int get() const { return v[dev()].load(std::memory_order_acquire); }
  void set(int x) { v[dev()].store(x, std::memory_order_release); }
};
struct PerDeviceFlag : PerDeviceInt {
  bool operator!() const { return get() == 0; }
  PerDeviceFlag& operator=(bool b) { set(b ? 1 : 0); return *this; }
};

// number of kernels this library has launched (bench.py reports it as gpu_launches)
extern unsigned long long g_launch_count;
#define GQ_LAUNCHED() (++::gq::g_launch_count)

// programmatic dependent launch on the serial kernel chains of K2 / K3+K4 (gq_set_pdl; ptx.cuh pdl_wait)
extern int g_pdl;
// <<<grid, block, smem, st>>> with the programmatic-serialization attribute when `pdl` is set; the kernel MUST call pdl_wait()
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl,
                                Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- K1
int syrk_upper(float* H, const void* X, int D, int T, int fmt, int mode, int variant, cudaStream_t st);
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_kind, uint64_t inner, uint64_t outer,
                 uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, int swizzle_atom32 = 0);

// general form: cu_dtype is a CUtensorMapDataType, cu_swizzle a CUtensorMapSwizzle
int make_tmap_f32_colgroups(CUtensorMap* out, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                            uint32_t box_groups);
int make_tmap_2d_raw(CUtensorMap* out, const void* base, int cu_dtype, uint64_t inner, uint64_t outer,
                     uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer, int cu_swizzle);

// ---- fp32 GEMM building block (SIMT):  C = beta*C + alpha * op(A) * B
enum : int {
  GEMM_TRANS_A = 1,   // A is stored (K x M) row-major  (C = A^T B); else (M x K)
  GEMM_C_UPPER = 2,   // only tiles that touch col >= row are computed (symmetric update)
  GEMM_A_UPPER = 4,   // A is upper triangular in its stored form -> shorter K range
  GEMM_B_UPPER = 8,   // B (K x N) is upper triangular: B[k][n] = 0 for k > n
};
struct GemmProblem {
  const float* A; const float* B; float* C;
  int M, N, K, lda, ldb, ldc;
};
constexpr int GEMM_MAX_BATCH = 48;
struct GemmBatch { GemmProblem p[GEMM_MAX_BATCH]; int count; };
int sgemm(const GemmProblem& pr, float alpha, float beta, int flags, cudaStream_t st);
int sgemm_batched(const GemmBatch& b, float alpha, float beta, int flags, cudaStream_t st);

// ---- fp32-faithful tensor-core GEMM  C += alpha * A^T B  (3 x TF32 split, tcgen05)
extern int g_w4a16_path;   // K6 dispatch: 0 auto; 1 never the mma.sync kernels (small / mid batch); 2 the weights-on-N tcgen05 kernel for batch <= 128; 3 auto without the mid-batch kernel; 4 the mid-batch kernel wherever it supports the shape
extern int g_syrk_kchunk;  // K1: 64-token k-blocks accumulated in tensor memory per partial sum (default 128 = 8192 tokens; 0 = a whole call)
extern int g_tf32_kchunk;  // 3xTF32 GEMM: 16-deep k-blocks accumulated in tensor memory per partial sum (default 8 = 128 deep; 0 = unlimited)
extern long long* g_solve_dbg;  // profiling hook: device buffer for the fused solver's clock64 stamps (first super-block of a call), or null
extern int g_solve_fused;  // K3/K4: 1 (default) = 128-column blocks run through the fused super-block kernel (near updates in tensor memory); 0 = one launch per block + near-update GEMMs
extern int g_solve_super;  // K4: blocks per super-block of the two-level lazy-batch update (1 = update everything after every block)
extern int g_use_tf32x3;   // 3 (default): K4 / Cholesky updates on tcgen05 3xTF32, truncation split (TF32_RAW_HI), 256 x 256 tiles on CTA pairs (cta_group::2) when there is more than one wave of tiles; 2: truncation split, single-CTA tiles only; 1: round-to-nearest split; 0: fp32 SIMT GEMM (A/B parity hook)
// a3 / b3_8 / b3_4: the same operands seen as (32 columns, rows, column groups), boxes of 16 rows x 4 / 8 / 4 groups: one TMA
// instruction per operand and k-block (groups_ok: both matrices have a multiple of 32 columns)
struct Tf32Maps { CUtensorMap a, b, c, a3, b3_8, b3_4; int groups_ok; };
// one GEMM of a batched launch: windows into the matrices the three tensor maps describe
struct Tf32Problem {
  int M, N, K;
  int a_col0, a_row0;      // A^T operand: element (k, m) at tensor-map coordinate (a_col0 + m, a_row0 + k)
  int b_col0, b_row0;
  int c_col0, c_row0;      // C element (m, n) at (c_col0 + n, c_row0 + m)
};
constexpr int TF32_MAX_BATCH = 56;
enum : int {
  TF32_UPPER_ONLY = 1,     // M == N: skip tiles entirely below the diagonal
  TF32_STORE = 2,          // C = alpha * A^T B (plain TMA store) instead of C += (TMA reduce-add)
  TF32_A_UPPER = 4,        // A (K x M) is upper triangular: k range of a tile ends at m0 + 128
  TF32_B_LOWER = 8,        // B (K x N) is lower triangular: k range of a tile starts at n0
  TF32_RAW_HI = 16,        // split stage keeps the landed fp32 tile as hi (the MMA truncates it) and writes only lo
  TF32_COLGROUPS = 32,     // (set by the launcher) operand windows start at multiples of 32 columns: 3D column-group boxes
};
int gemm_tf32x3_batch(const Tf32Maps& maps, const Tf32Problem* probs, int count, float alpha, int flags, cudaStream_t st);
int gemm_tf32x3_maps(Tf32Maps* maps, const float* A, int a_rows, int a_cols, int lda, const float* B, int b_rows,
                     int b_cols, int ldb, float* C, int c_rows, int c_cols, int ldc);
int gemm_tf32x3(const Tf32Maps& maps, int M, int N, int K, int a_row0, int a_col0, int b_row0, int b_col0, int c_row0,
                int c_col0, float alpha, int upper_only, cudaStream_t st);

// ---- K2
int hessian_gather(const float* H, int D, const int64_t* perm, float scale, float damp, const float* damp_dev, float damp_mult,
                   float* Hp, float* Arev, float* Hs, cudaStream_t st);
int chol_inverse_upper(float* A, int D, float* U, float* Wb, float* tmp, int* info_dev, cudaStream_t st);
size_t upper_packed_floats(int D);
int upper_tiles(float* H, float* packed, int D, int pack, cudaStream_t st);
int chol_factor_panel(float* A, int D, int o, int n, int rest, int* info_dev, cudaStream_t st);

// ---- K3/K4
int gptq_solve(float* Wt, const float* U, float* quant_t, void* codes_t, int codes_elem_size, float* err_t,
               float* scale_gr, float* qzero_gr, int R, int C, int group, float maxq, int mode, int a, int b,
               cudaStream_t st, const float* maxq_groups = nullptr);
int find_params(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, int sym, float* scale_gr,
                float* qzero_gr, cudaStream_t st);
int rtn_quantize(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* scale_gr,
                 const float* qzero_gr, float* quant_t, uint8_t* codes_t, cudaStream_t st);
int mse_refine(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* grid_dev,
               int n_grid, float norm, float* scale_gr, const float* qzero_gr, int* index_gr, cudaStream_t st);

// ---- K5/K6
int pack_int4(const uint8_t* codes, uint32_t* packed, int R, int C, cudaStream_t st);
int unpack_int4(const uint32_t* packed, uint8_t* codes, int R, int C, cudaStream_t st);
int transpose_pack_int4(const uint8_t* codes_t, uint8_t* codes, uint32_t* packed, int R, int C, cudaStream_t st);
constexpr int GQ_MAX_SEGMENTS = 8;   // linears that may be solved as one (row segments of the stacked weight)
int finalize_quant(const float* err_t, const float* quant_t, const void* orig_t, int fmt, int R, int C, const float* scale_gr,
                   const float* qzero_gr, int G, void* scale_out, uint8_t* qzero_out, const int64_t* perm, const float* U,
                   const int* seg_bounds, int nseg, double* stats, cudaStream_t st);
int gemm_f16_nt(const void* x, const void* w, void* y, int M, int N, int K, int fmt, cudaStream_t st);
// ---- f1: elementwise steps of the decoder block between the quantised linears (llama_ops.cu)
int llama_rms_norm(const void* x, const void* w, void* y, long long rows, int D, float eps, int fmt, cudaStream_t st);
int llama_rope(void* x, const void* cos_t, const void* sin_t, long long tokens, int heads, int hd, int seqlen, int fmt, cudaStream_t st);
int llama_silu_mul(const void* g, const void* u, void* out, long long n, int fmt, cudaStream_t st);
int reconstruct_dense(const uint8_t* q, const void* scale, const uint8_t* qzero, const int16_t* perm_inv, void* out, int R, int C,
                      int group, int fmt, cudaStream_t st);
int w4a16_dequant(const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, void* out, int N, int K, int group, int fmt,
                  cudaStream_t st);
size_t w4a16_workspace_bytes(int M, int N, int K);
size_t w4a16_gemv_workspace_bytes(int M, int K);
int w4a16_gemv(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm,
               void* y, void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
int w4a16_mma(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm,
              void* y, void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
bool w4a16_smallm_supports(int M, int N, int K, int group);
int w4a16_smallm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm,
                 void* y, void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
bool w4a16_midm_supports(int M, int N, int K, int group);
int w4a16_midm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm,
               void* y, void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
bool w4a16_wb_supports(int M, int N, int K, int group);
int w4a16_wb(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm, void* y,
             void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
int w4a16_gemm(const void* x, const uint32_t* qw, const void* scale_t, const uint8_t* qzero_t, const int32_t* perm,
               void* y, void* workspace, int M, int N, int K, int group, int fmt, cudaStream_t st);
int prep_weight_t(const void* W, int fmt, const int64_t* perm, float* Wt, void* orig_t, float* colsq, int R, int C,
                  cudaStream_t st);
int prep_weight_scatter(const void* W, int fmt, const int64_t* perm_inv, float* Wt, void* orig_t, float* colsq, int R, int C,
                        cudaStream_t st);
int solve_metrics(const float* err_t, const float* quant_t, const void* orig_t, int fmt, size_t n, double* out2, cudaStream_t st);

}  // namespace gq
