/* gptq_b200 -- C ABI of the B200-native GPTQ quantisation hot path (libgptq_b200.so).
 *
 * Drop-in boundary for 4PiR2/gptq-v2's native extension `gptq_c`
 * (reference: package/gptq/pybind.cpp:31-35 and the C-linkage launchers declared in
 * package/gptq/accumulate_hessian.cpp:4-17 and package/gptq/gptq_cpp.cpp:7-30).
 * Plain pointers and sizes only; every pointer is a DEVICE pointer owned by the caller,
 * borrowed for the duration of the call; results are written in place; nothing is
 * allocated.  `stream` is a cudaStream_t passed as void* (NULL = legacy default stream,
 * which is what the reference's <<<grid, block>>> launches use).  Kernels are launched
 * on the CURRENT device (the reference requires torch.cuda.set_device first, gptq_py.py:127).
 * Every function returns 0 on success or a GQ_ERR_* code; gq_error_string() names it.
 * The library is built for sm_100a only and has no CPU path.
 */
#ifndef GPTQ_B200_H_
#define GPTQ_B200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GQ_ABI_VERSION 1
#define GQ_DTYPE_F16 0
#define GQ_DTYPE_BF16 1
#define GQ_DTYPE_F32 2

int gq_abi_version(void);
const char* gq_error_string(int code);
/* number of kernels launched by this library since it was loaded (monotonic) */
unsigned long long gq_launch_count(void);

/* ---- K1: Hessian accumulation  H(fp32, D x D) += X^T X,  X (T x D) row-major fp16/bf16.
 * Replaces accumulate_hessian_{fp16,bf16}_fp32_cuda (accumulate_hessian.cpp:4-17,
 * accumulate_hessian_kernel.cu:67-84) and the Triton launcher
 * (gptq_triton/accumulate_hessian.py:134-149).
 * mode 0: only the upper triangle (col >= row) of H is updated
 * mode 1: upper triangle updated and mirrored into the lower one (Triton semantics)
 * mode 2: both triangles read-modify-written (CUTLASS semantics; H may be asymmetric)
 * Preconditions: D % 8 == 0, pointers 16-byte aligned. */
int gq_accumulate_hessian(void* H, const void* X, int size_hidden, int size_batch, int dtype, int mode, void* stream);
/* The reference's own launcher names and signatures (mode 2, default stream). */
int accumulate_hessian_fp16_fp32_cuda(void* H, const void* X, const int size_hidden, const int size_batch);
int accumulate_hessian_bf16_fp32_cuda(void* H, const void* X, const int size_hidden, const int size_batch);
/* test hook: variant 1 = single-CTA tcgen05 tiles, 2 = CTA-pair (cta_group::2) tiles */
int gq_accumulate_hessian_variant(void* H, const void* X, int size_hidden, int size_batch, int dtype, int mode,
                                  int variant, void* stream);

/* ---- K2: HessianHook.invert (gptq_py.py:39-66) after the host has chosen perm and damp.
 * Reads the upper triangle of H; forms Hp[i][j] = scale*H[perm i][perm j] (+damp, dead
 * diagonal -> 1) and U = chol_upper(Hp^-1).  perm may be NULL (identity).  Hp may be NULL.
 * work: gq_invert_workspace_floats(D) floats.  *info_dev (device int, caller zeroes it) is
 * set to 1 if a pivot is not positive (the caller adds damping and retries, as the
 * reference's NaN loop does).  Precondition: D % 4 == 0. */
size_t gq_invert_workspace_floats(int D);
int gq_hessian_invert(const float* H, int D, const int64_t* perm, float scale, float damp, float* Hp, float* U,
                      float* work, int* info_dev, void* stream);
/* The same with the damping taken from device memory: Hp = scale * H[perm][:, perm] + damp_dev[0] * damp_mult * I.  The caller
 * computes damp_ratio * mean(diag) (gptq_py.py:52) on the device and passes the try count as damp_mult (cumulative damping,
 * gptq_py.py:58), so that no host synchronisation sits between K1 and the factorisation. */
int gq_hessian_invert_dev(const float* H, int D, const int64_t* perm, float scale, const float* damp_dev, float damp_mult,
                          float* Hp, float* U, float* work, int* info_dev, void* stream);

/* ---- multi-GPU reduction payload (no reference counterpart: the reference is single-GPU).
 * K1 mode 0 maintains only the upper triangle of H, so only that is summed over ranks: the upper 128 x 128 tiles of
 * H (D, D) are copied into a contiguous buffer of gq_upper_packed_floats(D) floats (tile (bi, bj >= bi) at index
 * bi*nb - bi*(bi-1)/2 + bj - bi, nb = ceil(D/128), edge tiles zero-padded), all-reduced by NCCL as ONE message of about
 * half the bytes of H, and copied back.  gq_unpack_upper leaves everything outside those tiles untouched.  D % 4 == 0. */
size_t gq_upper_packed_floats(int D);
int gq_pack_upper(const float* H, int D, float* packed, void* stream);
int gq_unpack_upper(const float* packed, int D, float* H, void* stream);

/* ---- K3+K4, compat: quantize_range (gptq_cpp.cpp:33-84) on raw pointers.
 * Layout is the reference's transposed one: weights/quant/error/out_q are (C, R) row-major
 * with R = number of weight rows ("columns" in the reference's naming), hessian_inv (C, C).
 * scale/qzero (R) are static for the range; dequant is rounded through fp16.  out_q may be
 * NULL (the reference skips it for meta tensors).  Preconditions: R % 4 == 0, a % 4 == 0,
 * (b - a) a multiple of 32 (checked before anything is written; GQ_ERR_SHAPE otherwise). */
int gq_quantize_range(float* quant, const float* scale, int16_t* out_q, const float* qzero, float maxq,
                      const float* hessian_inv, float* weights, float* error, int C, int R, int a, int b,
                      void* stream);

/* ---- K3+K4, native: the whole blocked solve of one linear in the fp32 semantics of
 * gptq_quant(gptq_use_kernel=False) (gptq_py.py:158-193): per-group scale/zero from the
 * error-compensated weights, fp32 dequant.  Wt (C, R) in/out; outputs quant_t (C, R) (may be
 * NULL), codes_t (C, R) uint8, err_t (C, R), scale_gr/qzero_gr (C/group, R).
 * group in {32, 64, 128} (= the GPTQ block size, as llama.py:155 sets it). */
int gq_gptq_solve(float* Wt, const float* U, float* quant_t, uint8_t* codes_t, float* err_t, float* scale_gr,
                  float* qzero_gr, int R, int C, int group, float maxq, void* stream);
/* The kernel-path semantics of gptq_quant(gptq_use_kernel=True) (gptq_py.py:125-157) for ALL groups of one linear:
 * equivalent to the reference's loop `for g: gptq_quantize_range(..., scale[g], ..., qzero[g], ..., g*group, (g+1)*group)`
 * (static per-group scale_gr/qzero_gr (C/group, R) INPUTS, fp16-rounded dequant, out_q int16), without a host round
 * trip per group.  group in {32, 64, 128}. */
int gq_gptq_solve_static(float* Wt, const float* U, float* quant_t, int16_t* codes_t, float* err_t, const float* scale_gr,
                         const float* qzero_gr, int R, int C, int group, float maxq, void* stream);
/* ---- The remaining Quantizer options (quant.py:89-165, gptq_py.py:101-193) ----
 * General form of the two solves above over columns [a, b) (multiples of group; group in {32, 64, 128}):
 *   mode 0: gq_gptq_solve's semantics (the kernel computes asymmetric min/max parameters, scale_gr/qzero_gr OUT)
 *   mode 3: gq_gptq_solve_static's semantics (scale_gr/qzero_gr IN, 16-bit-rounded dequant, codes_t int16)
 *   mode 4: mode 0's fp32 arithmetic with caller-supplied scale_gr/qzero_gr (IN): what gptq_use_kernel=False does
 *           with a symmetric / MSE-refined / EXL2 Quantizer (gptq_py.py:169-192)
 * codes_t is uint8 in modes 0 and 4.  scale_gr/qzero_gr are the full (C/group, R) arrays.  maxq_groups: HOST array
 * of C/group floats (2^bits - 1 per group, gptq_py.py group_bit_widths), or NULL to use maxq for every group. */
int gq_gptq_solve_ex(float* Wt, const float* U, float* quant_t, void* codes_t, float* err_t, float* scale_gr,
                     float* qzero_gr, int R, int C, int group, const float* maxq_groups, float maxq, int mode, int a,
                     int b, void* stream);
/* Quantizer.find_params (quant.py:89-118) for n_groups consecutive groups of the transposed weights: Wt points at row
 * c0 of the (C, R) matrix, scale_gr/qzero_gr (OUT) and maxq_groups_dev (DEVICE array) at group c0/group.
 * sym = 0: min/max grid; sym = 1: scale = absmax * 2 / maxq + 1e-12, zero = round((maxq + 1) / 2). */
int gq_find_params(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, int sym,
                   float* scale_gr, float* qzero_gr, void* stream);
/* Round to nearest with given parameters (gptq_py.py:101-124, the hessian_hook=None branch: Quantizer.quantize /
 * dequantize per group, no error propagation): n_groups groups of `group` columns starting at row c0 of the (C, R)
 * arrays; every pointer is at that row / group.  q = clamp(round(w / scale) + zero, 0, maxq) -> codes_t (uint8, may be
 * NULL), (q - zero) * scale -> quant_t (may be NULL).  Any group size; R % 4 == 0. */
int gq_rtn_quantize(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* scale_gr,
                    const float* qzero_gr, float* quant_t, uint8_t* codes_t, void* stream);
/* Quantizer.mse (quant.py:152-156, use_kernel=False): per row and group, scale *= grid[argmin_i sum_c |deq(q(x_c;
 * scale * grid[i])) - x_c|^norm] (first minimum).  grid_dev: n_grid shrink factors on the device (the reference's
 * 1 - arange(0, max_shrink, 1/n_grid)); index_gr (n_groups, R) int32 receives the chosen index (may be NULL).
 * group <= 1024. */
int gq_mse_refine(const float* Wt, int R, int group, int n_groups, const float* maxq_groups_dev, const float* grid_dev,
                  int n_grid, float norm, float* scale_gr, const float* qzero_gr, int32_t* index_gr, void* stream);
/* Wt[c][r] = (float) W[r][perm[c]]; W (R, C) of dtype GQ_DTYPE_*; perm may be NULL. */
int gq_prep_weight_t(const void* W, int dtype, const int64_t* perm, float* Wt, int R, int C, void* stream);
/* Same, plus (each may be NULL) orig_t (C, R): the same elements in W's own dtype, kept for the error metrics after the
 * solver has consumed Wt; colsq (C) fp32: sum_r W[r][perm[c]]^2 (for gptq_norm, gptq_py.py:198). */
int gq_prep_weight_t2(const void* W, int dtype, const int64_t* perm, float* Wt, void* orig_t, float* colsq, int R, int C,
                      void* stream);
/* Scatter form of gq_prep_weight_t2 for callers that hold the INVERSE permutation (HessianHook.perm_inv,
 * gptq_py.py:49): Wt[perm_inv[k]][r] = (float) W[r][k], the same result as gq_prep_weight_t2 with perm, but W is read
 * in its own order and every write is a full 128-byte line (the gather form reads one element per 32-byte sector).
 * perm_inv may be NULL (identity); C must be even; orig_t / colsq as above (each may be NULL). */
int gq_prep_weight_t3(const void* W, int dtype, const int64_t* perm_inv, float* Wt, void* orig_t, float* colsq, int R, int C,
                      void* stream);
/* Result metrics of one linear in one pass (gptq_py.py:196, 232-233): out2[0] = sum err_t^2, out2[1] = sum (quant_t -
 * orig_t)^2 over n = R*C elements (n % 4 == 0); out2 is two doubles on the device, overwritten. */
int gq_solve_metrics(const float* err_t, const float* quant_t, const void* orig_t, int dtype, size_t n, double* out2,
                     void* stream);
/* Result assembly behind the solver in one call (gptq_py.py:196-199, 216-218, 231-237; replaces five torch passes).
 * The R rows may belong to up to 8 linears that share U and were solved as one stacked weight: seg_bounds (HOST array of
 * nseg + 1 ints, multiples of 4, seg_bounds[0] = 0, seg_bounds[nseg] = R; NULL = one segment) gives their row ranges.
 * stats (nseg x 4 doubles on the device, overwritten) = per segment [sum err_t^2, sum (quant_t - orig_t)^2,
 * sum_j sum_r orig_t[j][r]^2 / U[p_j][p_j]^2 with p_j = perm[j] (the reference's gptq_norm numerator: W unpermuted against the
 * permuted-order diag(U), reproduced as is), sum T(w^2) with the squares rounded to the weight dtype];
 * scale_out (R, G) of `dtype` and qzero_out (R, G) uint8 = the transposes of the solver's (G, R) fp32 tables.
 * err_t NULL: quant_t is used in its place (round to nearest: the caller reports 0).  U NULL: the third sum is 0;
 * perm NULL = identity. */
int gq_finalize_quant(const float* err_t, const float* quant_t, const void* orig_t, int dtype, int R, int C,
                      const float* scale_gr, const float* qzero_gr, int G, void* scale_out, uint8_t* qzero_out,
                      const int64_t* perm, const float* U, const int* seg_bounds, int nseg, double* stats, void* stream);

/* ---- K5: int4 storage (new in this build; the reference stores one code per byte, gptq_py.py:217).
 * packed (R, C/8) uint32, column 8k+i in bits [4i, 4i+4) of word k.  C % 8 == 0. */
int gq_pack_int4(const uint8_t* codes, uint32_t* packed, int R, int C, void* stream);
int gq_unpack_int4(const uint32_t* packed, uint8_t* codes, int R, int C, void* stream);
/* codes_t (C, R) -> codes (R, C) (may be NULL) and packed (R, C/8) (may be NULL) */
int gq_transpose_pack_int4(const uint8_t* codes_t, uint8_t* codes, uint32_t* packed, int R, int C, void* stream);

/* Dense 16-bit reconstruction from the packed format (quant.py:222-237 semantics: dense[n][k] = T(float(q - z) * float(s)),
 * one rounding; permuted column order kept, i.e. no perm_inv gather): the large-batch route of W4A16Linear (one
 * dequantisation + a dense GEMM once the batch is compute-bound).  dense (N, K) of dtype GQ_DTYPE_F16 / BF16, 16-byte
 * aligned; K % 8 == 0, group % 8 == 0, K % group == 0. */
int gq_w4a16_dequant(const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t, void* dense, int N, int K, int group,
                     int dtype, void* stream);

/* Dense 16-bit reconstruction straight from quant_meta (replaces construct_matrix + the perm_inv gather of
 * reconstruct_nn_linear, /root/reference/quant.py:222-237, 262-279, as the layer-by-layer driver calls it at llama.py:172):
 * dense[r][c] = T(float(q[r][p] - z[r][p / group]) * float(s[r][p / group])), p = perm_inv[c] (perm_inv NULL: p = c).
 * qweight (R, C) one code per byte, 16-byte aligned; scale (R, C / group) of dtype; qzero (R, C / group); perm_inv (C) int16
 * as quant_meta stores it; uniform group size, C % 16 == 0, C <= 32767.  One launch, no host synchronisation. */
int gq_reconstruct_dense(const uint8_t* qweight, const void* scale, const uint8_t* qzero, const int16_t* perm_inv, void* dense,
                         int R, int C, int group, int dtype, void* stream);

/* ---- f1 (driver costs): the elementwise steps of a Llama decoder block between the quantised linears, one pass each, with the
 * roundings of the PyTorch expressions the reference's model code evaluates when quantize_llama replays the block
 * (/root/reference/llama.py:121-125, 176-189 call transformers' LlamaDecoderLayer): no reference kernel counterpart.
 * All tensors contiguous, 16-byte aligned, dtype GQ_DTYPE_F16 / BF16.
 *   rms_norm: y = weight * T(float(x) * rsqrt(mean(float(x)^2) + eps)), rows x D, D % 8 == 0, D <= 8192 (y may alias x)
 *   rope:     in place on x (tokens, heads, head_dim): x * cos + rotate_half(x) * sin, every operation rounded to T; cos / sin
 *             (seqlen, head_dim), position = token % seqlen; head_dim % 16 == 0
 *   silu_mul: out = T(T(silu(gate)) * up) over n elements, n % 8 == 0 (out may alias gate or up) */
int gq_llama_rms_norm(const void* x, const void* weight, void* y, long long rows, int D, float eps, int dtype, void* stream);
int gq_llama_rope(void* x, const void* cos_t, const void* sin_t, long long tokens, int heads, int head_dim, int seqlen, int dtype,
                  void* stream);
int gq_llama_silu_mul(const void* gate, const void* up, void* out, long long n, int dtype, void* stream);
/* Dense 16-bit GEMM  y (M, N) = x (M, K) * w (N, K)^T  (fp32 accumulation, tcgen05 CTA pairs): the GEMM of the large-batch
 * route of K6 (gq_w4a16_dequant once, then this) -- the reference's nn.Linear forward on the reconstructed weights
 * (quant.py:262-279, llama.py:172-177) without a library GEMM.  x, w, y row-major contiguous of the same GQ_DTYPE_F16 / BF16,
 * 16-byte aligned; K % 8 == 0, N % 2 == 0. */
int gq_gemm_f16_nt(const void* x, const void* w, void* y, int M, int N, int K, int dtype, void* stream);
/* ---- K6: y (M, N) = x[:, perm] (M, K) * dequant(qweight)^T, replacing reconstruct_nn_linear +
 * nn.Linear (quant.py:262-279, llama.py:172-177).  x, y, scale_t share `dtype` (f16/bf16).  scale_t and
 * qzero_t are GROUP-MAJOR: (K/group, N) 16-bit and (K/group, N) uint8 (the transposes of quant_meta's scale /
 * qzero).  perm (K) int32 or NULL.  K % group == 0, group a power of two >= 32.
 * workspace: gq_w4a16_workspace_bytes(M, N, K) bytes of device scratch, 256-byte aligned (the permuted activations
 * and, for k-split work units, an fp32 accumulator). */
size_t gq_w4a16_workspace_bytes(int M, int N, int K);
int gq_w4a16_gemm(const void* x, const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t,
                  const int32_t* perm, void* y, void* workspace, int M, int N, int K, int group, int dtype,
                  void* stream);
/* Allocation-free batch 1-2 variant of gq_w4a16_gemm.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, layouts and result as gq_w4a16_gemm's CUDA-core route (w4a16_gemv.cu) minus `workspace`:
 * the shuffled copy of the activations lives in a file-scope device array of 64 KB that every GPU loading the library
 * holds until exit.  Limits: M <= 2 and K <= 16384 (GQ_ERR_LIMIT beyond them, before anything is launched); K % group == 0,
 * group % 32 == 0.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); the array's
 * device address is resolved once per device with cudaGetSymbolAddress.  All calls share the array and it is rewritten in
 * full on every call (nothing has to start at zero), so calls that overlap on one GPU (different streams or threads)
 * race on it: serialise them on one stream. */
int gq_w4a16_gemv_static(const void* x, const uint32_t* qweight, const void* scale_t, const uint8_t* qzero_t,
                         const int32_t* perm, void* y, int M, int N, int K, int group, int dtype, void* stream);

/* ---- building block exposed for tests: C = beta*C + alpha*op(A)*B in true fp32.
 * flags: 1 = A stored (K x M); 2 = only upper tiles of C; 4 = A upper-triangular; 8 = B upper-triangular */
int gq_sgemm(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, float alpha,
             float beta, int flags, void* stream);

/* ---- fp32-faithful tensor-core GEMM used by K4 and the Cholesky update:  C += alpha * A^T B  with A stored
 * (K x M), B (K x N), C (M x N), fp32, computed as 3 x tcgen05 kind::tf32 on hi/lo splits.  K % 16 == 0,
 * lda/ldb/ldc % 4 == 0.  upper_only: bit 0 = skip tiles entirely below the diagonal; bit 1 = C = alpha * A^T B
 * (plain store instead of accumulate; M % 32 == 0 then). */
int gq_gemm_tf32x3(const float* A, const float* B, float* C, int M, int N, int K, int lda, int ldb, int ldc, float alpha,
                   int upper_only, void* stream);
/* A/B parity hook: 0 routes K4 and the Cholesky trailing update through the fp32 SIMT GEMM instead of the
 * 3xTF32 tensor-core GEMM; 1 = 3xTF32 with x = rna_tf32(x) + lo; 2 = 3xTF32 with x = trunc_tf32(x) + lo (the landed
 * fp32 tile is the hi operand as it is: kind::tf32 reads only the upper 19 bits); 3 (default) = 2 plus 256 x 256 tiles on
 * CTA pairs (tcgen05 cta_group::2, each CTA stages half of B) when a launch has more than one wave of tiles.  Returns the
 * previous setting. */
int gq_set_tf32x3(int enable);
/* Tuning hook: blocks per super-block of K4's two-level lazy-batch update (default 8; 1 = the reference's schedule, one
 * full trailing update per block, gptq_py.py:139).  Returns the previous setting. */
int gq_set_solve_super(int blocks);
/* Accuracy / speed hook of K1: token range (in 64-token k-blocks) accumulated in tensor memory before the partial sum is added
 * to H in fp32 round-to-nearest (default 128 = 8192 tokens; 0 = the whole call, round 1's behaviour).  Returns the old value. */
int gq_set_syrk_kchunk(int kblocks);
/* Accuracy / speed hook of the 3xTF32 GEMM: the tensor core adds every MMA into its fp32 accumulator with truncation, so a
 * k range is accumulated in tensor memory in chunks of `kblocks` x 16 and each chunk's partial sum is added to C by a
 * round-to-nearest reduce-add (default 8 = 128-deep chunks; 0 = the whole range at once, round 1's behaviour).
 * Returns the previous setting. */
int gq_set_tf32_kchunk(int kblocks);
/* Profiling hook: a device buffer of >= 256 int64 that receives clock64() stamps of the first chain warp of CTA 0 for the first
 * super-block of every following fused solve (tools/solve_timeline.py decodes them); NULL switches it off.  Returns the old pointer. */
void* gq_set_solve_debug(void* dev_buffer);
/* A/B hook: 1 (default) = 128-column blocks are solved by the fused super-block kernel (one launch per up to four blocks, their
 * mutual updates accumulated in tensor memory); 0 = one launch per block plus near-update GEMM launches (round 1's schedule).
 * The environment variable GQ_SOLVE_FUSED=0 selects 0 at load time.  Returns the previous setting. */
int gq_set_solve_fused(int on);
/* Programmatic dependent launch on the serial kernel chains (Cholesky panel <-> trailing update, solver super-block <-> far
 * update): 1 (default) lets a kernel's prologue overlap the tail of its predecessor, 0 restores plain stream order (A/B
 * hook; results are identical).  Returns the previous value. */
int gq_set_pdl(int on);
/* A/B hook for K6's dispatch: 0 = automatic (batch <= 8: warp-level mma kernel; batch 9..16: its windowed form; otherwise tcgen05 with the weights on the M
 * side through tensor memory, 64 or 256 batch columns per pass), 1 = never the small-batch kernel, 2 = batch <= 128 through
 * the experimental tcgen05 kernel with the weights on the N side (w4a16_wb.cu; correct, not yet faster), 3 = automatic
 * without the batch 9..16 kernel (w4a16_midm.cu), i.e. the dispatch before it existed, 4 = that kernel for every batch <= 16
 * it supports.  Returns the previous setting. */
int gq_set_w4a16_path(int path);

#ifdef __cplusplus
}
#endif
#endif /* GPTQ_B200_H_ */
