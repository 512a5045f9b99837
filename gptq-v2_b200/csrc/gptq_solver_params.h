#pragma once
namespace gq {
struct SolveParams {
  float* Wt;            // (C, R) in/out
  const float* U;       // (C, C)
  float* quant_t;       // (C, R) out, may be null
  void* codes_t;        // (C, R) out (uint8 or int16), may be null
  float* err_t;         // (C, R) out
  float* scale;         // MODE 0: out (R) for this group; MODE 1: in (R)
  float* qzero;         // same
  int R, C;
  int c0;               // first column of the block
  float maxq;
  int static_params;    // MODE 0 only: scale/qzero are inputs (symmetric / MSE-refined / EXL2 parameters)
};
}  // namespace gq
