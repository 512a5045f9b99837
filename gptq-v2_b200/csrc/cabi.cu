// extern "C" surface of libgptq_b200.so -- see include/gptq_b200.h for the contract.
#include "../../include/gptq_b200.h"
#include "gptq_b200_internal.h"

using namespace gq;

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

namespace gq { unsigned long long g_launch_count = 0; int g_use_tf32x3 = 3; int g_solve_super = 8; int g_solve_fused = 1; int g_tf32_kchunk = 8; int g_syrk_kchunk = 128; long long* g_solve_dbg = nullptr; int g_w4a16_path = 0; int g_pdl = 1; }

extern "C" {

unsigned long long gq_launch_count(void) { return gq::g_launch_count; }

int gq_abi_version(void) { return GQ_ABI_VERSION; }

const char* gq_error_string(int code) {
  switch (code) {
    case GQ_OK: return "ok";
    case GQ_ERR_CUDA: return "CUDA runtime error (launch or API call failed)";
    case GQ_ERR_SHAPE: return "unsupported shape (see preconditions in gptq_b200.h)";
    case GQ_ERR_DTYPE: return "unsupported dtype";
    case GQ_ERR_ALIGN: return "pointer is not sufficiently aligned";
    case GQ_ERR_DRIVER: return "cuTensorMapEncodeTiled driver entry point not found";
    case GQ_ERR_TENSORMAP: return "cuTensorMapEncodeTiled rejected the tensor map";
    case GQ_ERR_NOT_SPD: return "matrix is not positive definite";
    case GQ_ERR_ARG: return "invalid argument";
    case GQ_ERR_LIMIT: return "size exceeds the fixed scratch of this allocation-free entry point (see gptq_b200.h)";
    default: return "unknown error";
  }
}

int gq_accumulate_hessian_variant(void* H, const void* X, int size_hidden, int size_batch, int dtype, int mode,
                                  int variant, void* stream) {
  if (mode < 0 || mode > 2) return GQ_ERR_ARG;
  return syrk_upper(static_cast<float*>(H), X, size_hidden, size_batch, dtype, mode, variant, S(stream));
}
int gq_accumulate_hessian(void* H, const void* X, int size_hidden, int size_batch, int dtype, int mode, void* stream) {
  return gq_accumulate_hessian_variant(H, X, size_hidden, size_batch, dtype, mode, 2, stream);
}
int accumulate_hessian_fp16_fp32_cuda(void* H, const void* X, const int size_hidden, const int size_batch) {
  return gq_accumulate_hessian(H, X, size_hidden, size_batch, GQ_DTYPE_F16, 2, nullptr);
}
int accumulate_hessian_bf16_fp32_cuda(void* H, const void* X, const int size_hidden, const int size_batch) {
  return gq_accumulate_hessian(H, X, size_hidden, size_batch, GQ_DTYPE_BF16, 2, nullptr);
}

size_t gq_invert_workspace_floats(int D) {
  const size_t d = (size_t)D;
  return 2 * d * d + d * d / 2 + 32 * d + 64;     // A (reversed matrix / factor / V), W = V^T, T^T blocks
}

static int hessian_invert_impl(const float* H, int D, const int64_t* perm, float scale, float damp, const float* damp_dev,
                               float damp_mult, float* Hp, float* U, float* work, int* info_dev, void* stream) {
  if (D <= 0) return GQ_ERR_ARG;
  if (D % 4 != 0) return GQ_ERR_SHAPE;
  float* A = work;
  float* Wb = work + (size_t)D * D;
  float* tmp = work + 2 * (size_t)D * D;
  int rc = hessian_gather(H, D, perm, scale, damp, damp_dev, damp_mult, Hp, A, Wb, S(stream));   // Wb doubles as the symmetric copy
  if (rc != GQ_OK) return rc;
  return chol_inverse_upper(A, D, U, Wb, tmp, info_dev, S(stream));
}

int gq_hessian_invert(const float* H, int D, const int64_t* perm, float scale, float damp, float* Hp, float* U,
                      float* work, int* info_dev, void* stream) {
  return hessian_invert_impl(H, D, perm, scale, damp, nullptr, 0.f, Hp, U, work, info_dev, stream);
}

int gq_hessian_invert_dev(const float* H, int D, const int64_t* perm, float scale, const float* damp_dev, float damp_mult,
                          float* Hp, float* U, float* work, int* info_dev, void* stream) {
  if (!damp_dev) return GQ_ERR_ARG;
  return hessian_invert_impl(H, D, perm, scale, 0.f, damp_dev, damp_mult, Hp, U, work, info_dev, stream);
}

size_t gq_upper_packed_floats(int D) { return D > 0 ? upper_packed_floats(D) : 0; }
int gq_pack_upper(const float* H, int D, float* packed, void* stream) {
  return upper_tiles(const_cast<float*>(H), packed, D, 1, S(stream));
}
int gq_unpack_upper(const float* packed, int D, float* H, void* stream) {
  return upper_tiles(H, const_cast<float*>(packed), D, 0, S(stream));
}

int gq_quantize_range(float* quant, const float* scale, int16_t* out_q, const float* qzero, float maxq,
                      const float* hessian_inv, float* weights, float* error, int C, int R, int a, int b,
                      void* stream) {
  return gptq_solve(weights, hessian_inv, quant, out_q, 2, error, const_cast<float*>(scale),
                    const_cast<float*>(qzero), R, C, 128, maxq, 1, a, b, S(stream));
}

int gq_gptq_solve(float* Wt, const float* U, float* quant_t, uint8_t* codes_t, float* err_t, float* scale_gr,
                  float* qzero_gr, int R, int C, int group, float maxq, void* stream) {
  return gptq_solve(Wt, U, quant_t, codes_t, 1, err_t, scale_gr, qzero_gr, R, C, group, maxq, 0, 0, C, S(stream));
}

int gq_gptq_solve_static(float* Wt, const float* U, float* quant_t, int16_t* codes_t, float* err_t, const float* scale_gr,
                         const float* qzero_gr, int R, int C, int group, float maxq, void* stream) {
  return gptq_solve(Wt, U, quant_t, codes_t, 2, err_t, const_cast<float*>(scale_gr), const_cast<float*>(qzero_gr), R, C,
                    group, maxq, 3, 0, C, S(stream));
}

int gq_gptq_solve_ex(float* Wt, const float* U, float* quant_t, void* codes_t, float* err_t, float* scale_gr,
                     float* qzero_gr, int R, int C, int group, const float* maxq_groups, float maxq, int mode, int a,
                     int b, void* stream) {
  if (mode != 0 && mode != 3 && mode != 4) return GQ_ERR_ARG;
  return gptq_solve(Wt, U, quant_t, codes_t, mode == 3 ? 2 : 1, err_t, scale_gr, qzero_gr, R, C, group, maxq, mode, a, b,
                    S(stream), maxq_groups);
}

int gq_find_params(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, int sym,
                   float* scale_gr, float* qzero_gr, void* stream) {
  return find_params(Wt, R, group, n_groups, maxq_groups_dev, sym, scale_gr, qzero_gr, S(stream));
}

int gq_rtn_quantize(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* scale_gr,
                    const float* qzero_gr, float* quant_t, uint8_t* codes_t, void* stream) {
  return rtn_quantize(Wt, R, group, n_groups, maxq_groups_dev, scale_gr, qzero_gr, quant_t, codes_t, S(stream));
}

int gq_mse_refine(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* grid_dev,
                  int n_grid, float norm, float* scale_gr, const float* qzero_gr, int32_t* index_gr, void* stream) {
  return mse_refine(Wt, R, group, n_groups, maxq_groups_dev, grid_dev, n_grid, norm, scale_gr, qzero_gr, index_gr,
                    S(stream));
}

int gq_prep_weight_t(const void* W, int dtype, const int64_t* perm, float* Wt, int R, int C, void* stream) {
  return prep_weight_t(W, dtype, perm, Wt, nullptr, nullptr, R, C, S(stream));
}
int gq_prep_weight_t2(const void* W, int dtype, const int64_t* perm, float* Wt, void* orig_t, float* colsq, int R, int C,
                      void* stream) {
  return prep_weight_t(W, dtype, perm, Wt, orig_t, colsq, R, C, S(stream));
}
int gq_prep_weight_t3(const void* W, int dtype, const int64_t* perm_inv, float* Wt, void* orig_t, float* colsq, int R, int C,
                      void* stream) {
  return prep_weight_scatter(W, dtype, perm_inv, Wt, orig_t, colsq, R, C, S(stream));
}
int gq_solve_metrics(const float* err_t, const float* quant_t, const void* orig_t, int dtype, size_t n, double* out2,
                     void* stream) {
  return solve_metrics(err_t, quant_t, orig_t, dtype, n, out2, S(stream));
}

int gq_finalize_quant(const float* err_t, const float* quant_t, const void* orig_t, int dtype, int R, int C,
                      const float* scale_gr, const float* qzero_gr, int G, void* scale_out, uint8_t* qzero_out,
                      const int64_t* perm, const float* U, const int* seg_bounds, int nseg, double* stats, void* stream) {
  return finalize_quant(err_t, quant_t, orig_t, dtype, R, C, scale_gr, qzero_gr, G, scale_out, qzero_out, perm, U, seg_bounds,
                        nseg, stats, S(stream));
}

int gq_pack_int4(const uint8_t* codes, uint32_t* packed, int R, int C, void* stream) {
  return pack_int4(codes, packed, R, C, S(stream));
}
int gq_unpack_int4(const uint32_t* packed, uint8_t* codes, int R, int C, void* stream) {
  return unpack_int4(packed, codes, R, C, S(stream));
}
int gq_transpose_pack_int4(const uint8_t* codes_t, uint8_t* codes, uint32_t* packed, int R, int C, void* stream) {
  return transpose_pack_int4(codes_t, codes, packed, R, C, S(stream));
}

int gq_llama_rms_norm(const void* x, const void* weight, void* y, long long rows, int D, float eps, int dtype, void* stream) {
  return llama_rms_norm(x, weight, y, rows, D, eps, dtype, S(stream));
}
int gq_llama_rope(void* x, const void* cos_t, const void* sin_t, long long tokens, int heads, int head_dim, int seqlen, int dtype,
                  void* stream) {
  return llama_rope(x, cos_t, sin_t, tokens, heads, head_dim, seqlen, dtype, S(stream));
}
int gq_llama_silu_mul(const void* gate, const void* up, void* out, long long n, int dtype, void* stream) {
  return llama_silu_mul(gate, up, out, n, dtype, S(stream));
}
int gq_reconstruct_dense(const uint8_t* qweight, const void* scale, const uint8_t* qzero, const int16_t* perm_inv, void* dense,
                         int R, int C, int group, int dtype, void* stream) {
  return reconstruct_dense(qweight, scale, qzero, perm_inv, dense, R, C, group, dtype, S(stream));
}
int gq_w4a16_dequant(const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t, void* dense, int N, int K, int group,
                     int dtype, void* stream) {
  return w4a16_dequant(qweight, scale_t, qzero_t, dense, N, K, group, dtype, S(stream));
}
int gq_gemm_f16_nt(const void* x, const void* w, void* y, int M, int N, int K, int dtype, void* stream) {
  return gemm_f16_nt(x, w, y, M, N, K, dtype, S(stream));
}
size_t gq_w4a16_workspace_bytes(int M, int N, int K) { return w4a16_workspace_bytes(M, N, K); }
int gq_w4a16_gemm(const void* x, const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t,
                  const int32_t* perm, void* y, void* workspace, int M, int N, int K, int group, int dtype,
                  void* stream) {
  return w4a16_gemm(x, qweight, scale_t, qzero_t, perm, y, workspace, M, N, K, group, dtype, S(stream));
}

int gq_sgemm(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, float alpha,
             float beta, int flags, void* stream) {
  GemmProblem p{A, B, C, M, N, K, lda, ldb, ldc};
  return sgemm(p, alpha, beta, flags, S(stream));
}

int gq_set_solve_super(int blocks) {
  int old = gq::g_solve_super;
  if (blocks >= 1 && blocks <= 16) gq::g_solve_super = blocks;
  return old;
}
int gq_set_syrk_kchunk(int kblocks) {
  int old = gq::g_syrk_kchunk;
  if (kblocks >= 0 && kblocks <= (1 << 20)) gq::g_syrk_kchunk = kblocks;
  return old;
}
int gq_set_tf32_kchunk(int kblocks) {
  int old = gq::g_tf32_kchunk;
  if (kblocks >= 0 && kblocks <= 4096) gq::g_tf32_kchunk = kblocks;
  return old;
}
void* gq_set_solve_debug(void* dev_buffer) {
  void* old = gq::g_solve_dbg;
  gq::g_solve_dbg = static_cast<long long*>(dev_buffer);
  return old;
}
int gq_set_solve_fused(int on) {
  int old = gq::g_solve_fused;
  gq::g_solve_fused = on ? 1 : 0;
  return old;
}
int gq_set_pdl(int on) {
  int old = gq::g_pdl;
  if (on == 0 || on == 1) gq::g_pdl = on;
  return old;
}
int gq_set_w4a16_path(int path) {
  int old = gq::g_w4a16_path;
  if (path >= 0 && path <= 4) gq::g_w4a16_path = path;
  return old;
}
int gq_set_tf32x3(int enable) { int old = gq::g_use_tf32x3; gq::g_use_tf32x3 = (enable >= 0 && enable <= 3) ? enable : 1; return old; }

int gq_gemm_tf32x3(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, float alpha,
                   int upper_only, void* stream) {
  Tf32Maps maps;
  int rc = gemm_tf32x3_maps(&maps, A, K, M, lda, B, K, N, ldb, C, M, N, ldc);
  if (rc != GQ_OK) return rc;
  Tf32Problem q{M, N, K, 0, 0, 0, 0, 0, 0};
  return gemm_tf32x3_batch(maps, &q, 1, alpha, upper_only & (TF32_UPPER_ONLY | TF32_STORE), S(stream));
}

}  // extern "C"
