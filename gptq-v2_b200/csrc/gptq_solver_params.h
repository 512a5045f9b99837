#pragma once
#include <cuda.h>
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
constexpr int SUP_MAXBLK = 4;   // blocks per fused super-block launch: 3 accumulator regions of 128 TMEM columns
struct SuperParams {
  float* Wt;            // (C, R) in (and out in MODE 1)
  const float* U;       // (C, C)
  float* quant_t;       // (C, R) out, may be null
  void* codes_t;        // (C, R) out, may be null
  float* err_t;         // (C, R) out
  float* scale;         // parameters of the first block; block b at scale + b * param_stride
  float* qzero;
  int param_stride;     // R for per-group parameters, 0 when one (R) vector serves the whole range (quantize_range)
  int R, C;
  int c0;               // first column of the super-block
  int nblk;             // 1..SUP_MAXBLK blocks of 128 columns
  float maxq[SUP_MAXBLK];
  int static_params;    // MODE 0 only: scale / qzero are inputs
  long long* dbg;       // profiling hook (gq_set_solve_debug): clock64 stamps of CTA 0's first chain warp, or null
};

// U as (32 columns, rows, column groups): box of 8 rows x 4 / 8 / 12 column groups = the later columns of one / two / three
// blocks of a super-block in ONE TMA instruction per 8-deep k-slice (m[g] holds the box with 4 (g + 1) groups)
// upk[g]: plain box of 16 rows x (128 - 16 g) columns = rows [16 g, 16 g + 16) of a diagonal block of U from their first kept
// column on: the packed diagonal block arrives in 8 TMA instructions (128 row-wise bulk copies kept the TMA issue slot busy
// for ~7000 cycles per block, in front of the ring loads)
struct SuperMaps { CUtensorMap m[3]; CUtensorMap upk[8]; };

}  // namespace gq
