// `gptq_c` -- the reference's native module (package/gptq/pybind.cpp:31-35) on top of libgptq_b200.so.
//
// Same module name, function names, argument lists and void returns as the reference's pybind module, so that the
// reference's own package/gptq/__init__.py (`import gptq_c`) works unchanged with this build:
//     accumulate_hessian_fp16_fp32(Tensor& H, const Tensor& X)        package/gptq/accumulate_hessian.cpp:20-31
//     accumulate_hessian_bf16_fp32(Tensor& H, const Tensor& X)        package/gptq/accumulate_hessian.cpp:34-45
//     gptq_quantize_range(quant, scale, out_q, qzero, maxq, hessian_inv, weights, error, a, b)   package/gptq/gptq_cpp.cpp:33-84
// Differences, all on the safe side: kernels launch on torch's CURRENT stream of the tensors' device
// (the reference uses the legacy default stream of the current device), a failing launch raises instead of being
// ignored (accumulate_hessian.cpp:25-30 drops the CUTLASS status), and shapes / dtypes / contiguity are checked.
// No arithmetic lives here: every call forwards raw device pointers through the C ABI of include/gptq_b200.h.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include "../../include/gptq_b200.h"

namespace {

void check_rc(int rc, const char* what) {
  TORCH_CHECK(rc == 0, what, ": ", gq_error_string(rc), " (code ", rc, ")");
}

void accumulate(torch::Tensor& H, const torch::Tensor& X, int dtype, c10::ScalarType st) {
  TORCH_CHECK(H.is_cuda() && X.is_cuda() && H.device() == X.device(), "gptq_c: CUDA tensors on one device expected");
  TORCH_CHECK(H.scalar_type() == torch::kFloat32 && X.scalar_type() == st, "gptq_c: dtype mismatch");
  TORCH_CHECK(H.is_contiguous() && X.is_contiguous(), "gptq_c: contiguous (row-major) tensors expected");
  TORCH_CHECK(H.dim() == 2 && X.dim() == 2 && H.size(0) == H.size(1) && H.size(1) == X.size(1), "H shape mismatch");
  const c10::cuda::CUDAGuard guard(H.device());
  void* stream = at::cuda::getCurrentCUDAStream(H.device().index()).stream();
  // mode 2: both triangles read-modify-written, the CUTLASS kernel's semantics (accumulate_hessian_kernel.cu:47-52)
  check_rc(gq_accumulate_hessian(H.data_ptr(), X.data_ptr(), (int)X.size(1), (int)X.size(0), dtype, 2, stream),
           "accumulate_hessian");
}

// Internal linkage on purpose: the reference's own extension defines global functions with these very names; if both
// libraries are loaded into one process (the tests do that) the dynamic linker would otherwise bind this module's calls to
// whichever definition was loaded first.
void accumulate_hessian_fp16_fp32(torch::Tensor& H, const torch::Tensor& X) { accumulate(H, X, GQ_DTYPE_F16, torch::kFloat16); }
void accumulate_hessian_bf16_fp32(torch::Tensor& H, const torch::Tensor& X) { accumulate(H, X, GQ_DTYPE_BF16, torch::kBFloat16); }

void quantize_range(torch::Tensor quant, torch::Tensor scale, torch::Tensor out_q, torch::Tensor qzero, float maxq,
                    torch::Tensor hessian_inv, torch::Tensor weights, torch::Tensor error, int a, int b) {
  TORCH_CHECK(weights.dim() == 2 && hessian_inv.dim() == 2 && weights.size(0) == hessian_inv.size(1), "H shape mismatch");   // gptq_cpp.cpp:49
  const bool want_q = !out_q.is_meta();                                                                                 // gptq_cpp.cpp:58
  for (const torch::Tensor* t : {&quant, &scale, &qzero, &hessian_inv, &weights, &error}) {
    TORCH_CHECK(t->is_cuda() && t->device() == weights.device(), "gptq_c: CUDA tensors on one device expected");
    TORCH_CHECK(t->scalar_type() == torch::kFloat32 && t->is_contiguous(), "gptq_c: contiguous float32 tensors expected");
  }
  if (want_q)
    TORCH_CHECK(out_q.is_cuda() && out_q.scalar_type() == torch::kInt16 && out_q.is_contiguous() && out_q.sizes() == weights.sizes(),
                "gptq_c: out_q must be a contiguous int16 CUDA tensor shaped like weights (or a meta tensor)");
  const int C = (int)weights.size(0), R = (int)weights.size(1);
  TORCH_CHECK(quant.sizes() == weights.sizes() && error.sizes() == weights.sizes() && scale.numel() == R && qzero.numel() == R,
              "gptq_c: quant / error (C, R) and scale / qzero (R) expected");
  const c10::cuda::CUDAGuard guard(weights.device());
  void* stream = at::cuda::getCurrentCUDAStream(weights.device().index()).stream();
  check_rc(gq_quantize_range(quant.data_ptr<float>(), scale.data_ptr<float>(), want_q ? out_q.data_ptr<int16_t>() : nullptr,
                             qzero.data_ptr<float>(), maxq, hessian_inv.data_ptr<float>(), weights.data_ptr<float>(),
                             error.data_ptr<float>(), C, R, a, b, stream),
           "gptq_quantize_range");
}

}  // namespace

PYBIND11_MODULE(gptq_c, m) {
  m.def("accumulate_hessian_fp16_fp32", &accumulate_hessian_fp16_fp32, "H += X.t() @ X (FP16 x FP16 = FP32 matmul)");
  m.def("accumulate_hessian_bf16_fp32", &accumulate_hessian_bf16_fp32, "H += X.t() @ X (BF16 x BF16 = FP32 matmul)");
  m.def("gptq_quantize_range", &quantize_range, "Quantize Weights within Range Using GPTQ");
}
