"""Drop-in for the reference's `gptq` extension package (package/gptq/__init__.py:6-28):

    from gptq import accumulate_hessian, gptq_quantize_range        # gptq_py.py:3

Same names, argument order, in-place semantics and dtype dispatch, backed by the
sm_100a library libgptq_b200.so instead of `gptq_c`.  The additional ops below
(hessian_syrk, hessian_invert, gptq_solve, pack_int4, unpack_int4, w4a16_gemm)
are what this build's own gptq_py.HessianHook / gptq_quant call.
"""
from __future__ import annotations

import ctypes

import torch

from . import _cabi
from ._cabi import GptqB200Error, check, lib, ptr, require_contig, require_cuda, stream_of

__all__ = ['accumulate_hessian', 'gptq_quantize_range', 'hessian_syrk', 'hessian_invert', 'w4a16_dequant', 'reconstruct_dense', 'llama_rms_norm', 'llama_rope_', 'llama_silu_mul', 'pack_upper', 'unpack_upper', 'prep_weight_t', 'prep_weight_t2', 'solve_metrics', 'finalize_quant', 'gptq_solve_ex', 'find_params', 'mse_refine', 'find_params_cols', 'mse_refine_cols', 'rtn_quantize',
           'gptq_solve', 'gptq_solve_static', 'pack_int4', 'unpack_int4', 'transpose_pack_int4', 'w4a16_gemm', 'w4a16_gemv_static', 'gemm_f16_nt', 'sgemm', 'gemm_tf32x3', 'set_tf32x3', 'set_solve_super', 'set_solve_fused', 'set_tf32_kchunk', 'set_syrk_kchunk', 'set_w4a16_path', 'set_pdl',
           'GptqB200Error']


# ----------------------------------------------------------------------------
# reference surface
# ----------------------------------------------------------------------------
def accumulate_hessian(mat_hessian: torch.Tensor, mat_input: torch.Tensor) -> None:
    """H += X^T X (fp16 or bf16 X, fp32 H), in place; both triangles are updated
    (the CUTLASS kernel's semantics, package/gptq/accumulate_hessian_kernel.cu:47-52)."""
    if mat_input.dtype not in (torch.float16, torch.bfloat16):
        raise NotImplementedError                       # package/gptq/__init__.py:12-13
    hessian_syrk(mat_hessian, mat_input, mode=2)


def gptq_quantize_range(quant: torch.Tensor, scale: torch.Tensor, out_q: torch.Tensor, qzero: torch.Tensor,
                        maxq: float, hessian_inv: torch.Tensor, weights: torch.Tensor, error: torch.Tensor,
                        a: int, b: int) -> None:
    """Columns [a, b) of the transposed (C, R) weight matrix: quantise, dequantise through
    fp16, propagate the error inside the range and to every later column
    (package/gptq/gptq_cpp.cpp:33-84).  All tensors are updated in place."""
    if out_q is not None and out_q.device.type == 'meta':
        out_q = None                                     # gptq_cpp.cpp:58
    dev = require_cuda(quant, scale, out_q, qzero, hessian_inv, weights, error)
    require_contig(quant=quant, scale=scale, out_q=out_q, qzero=qzero, hessian_inv=hessian_inv, weights=weights,
                   error=error)
    C_, R_ = weights.shape
    if hessian_inv.size(1) != C_:
        raise RuntimeError('H shape mismatch')           # gptq_cpp.cpp:49
    for t in (quant, scale, qzero, hessian_inv, weights, error):
        if t.dtype != torch.float32:
            raise GptqB200Error('gptq_quantize_range expects float32 tensors')
    if out_q is not None and out_q.dtype != torch.int16:
        raise GptqB200Error('out_q must be int16')
    with torch.cuda.device(dev):
        check(lib.gq_quantize_range(ptr(quant), ptr(scale), ptr(out_q), ptr(qzero), float(maxq), ptr(hessian_inv),
                                    ptr(weights), ptr(error), C_, R_, int(a), int(b), stream_of(weights)),
              'gq_quantize_range')


# ----------------------------------------------------------------------------
# native ops
# ----------------------------------------------------------------------------
def hessian_syrk(H: torch.Tensor, X: torch.Tensor, mode: int = 1, variant: int = 2) -> None:
    """K1.  H (D, D) fp32 += X^T X with X (T, D) fp16/bf16.
    mode 0 upper triangle only, 1 upper + mirrored lower, 2 both triangles RMW."""
    dev = require_cuda(H, X)
    require_contig(H=H, X=X)
    if X.dim() != 2 or H.dim() != 2 or H.shape[0] != H.shape[1] or H.shape[1] != X.shape[1]:
        raise GptqB200Error(f'shape mismatch: H {tuple(H.shape)} X {tuple(X.shape)}')
    if H.dtype != torch.float32 or X.dtype not in (torch.float16, torch.bfloat16):
        raise NotImplementedError
    T, D = X.shape
    with torch.cuda.device(dev):
        check(lib.gq_accumulate_hessian_variant(ptr(H), ptr(X), D, T, _cabi.DTYPE_CODE[X.dtype], mode, variant,
                                                stream_of(H)), 'gq_accumulate_hessian')


def hessian_invert(H: torch.Tensor, perm, scale: float, damp, want_hp: bool = True, damp_mult: float = 1.0):
    """K2.  Returns (Hp, U, info) with Hp = scale*H[perm][:, perm] + damp*I and
    U = chol_upper(Hp^-1); info is a device int32 (1 = a pivot was not positive).
    damp: a Python float, or a 0-dim fp32 CUDA tensor (then the damping is damp * damp_mult, read on the device: no host
    synchronisation between K1 and the factorisation)."""
    dev = require_cuda(H, perm)
    require_contig(H=H, perm=perm)
    D = H.shape[0]
    if perm is not None and perm.dtype != torch.int64:
        raise GptqB200Error('perm must be int64')
    U = torch.empty_like(H)
    Hp = torch.empty_like(H) if want_hp else None
    work = torch.empty(lib.gq_invert_workspace_floats(D), dtype=torch.float32, device=dev)
    info = torch.zeros((), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        if isinstance(damp, torch.Tensor):
            if damp.dtype != torch.float32 or damp.numel() != 1 or damp.device != dev:
                raise GptqB200Error('damp tensor must be one fp32 value on the device of H')
            check(lib.gq_hessian_invert_dev(ptr(H), D, ptr(perm), float(scale), ptr(damp), float(damp_mult), ptr(Hp), ptr(U),
                                            ptr(work), ptr(info), stream_of(H)), 'gq_hessian_invert_dev')
        else:
            check(lib.gq_hessian_invert(ptr(H), D, ptr(perm), float(scale), float(damp) * float(damp_mult), ptr(Hp), ptr(U),
                                        ptr(work), ptr(info), stream_of(H)), 'gq_hessian_invert')
    return Hp, U, info


def pack_upper(H: torch.Tensor) -> torch.Tensor:
    """Upper 128 x 128 tiles of H (D, D) fp32 -> one contiguous fp32 buffer (the all-reduce payload of sharded.py)."""
    dev = require_cuda(H)
    require_contig(H=H)
    D = H.shape[0]
    assert H.dtype == torch.float32 and H.shape == (D, D)
    packed = torch.empty(lib.gq_upper_packed_floats(D), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_pack_upper(ptr(H), D, ptr(packed), stream_of(H)), 'gq_pack_upper')
    return packed


def unpack_upper(packed: torch.Tensor, H: torch.Tensor) -> None:
    """Inverse of pack_upper, in place: the upper tiles of H are overwritten, the rest is left alone."""
    dev = require_cuda(packed, H)
    require_contig(packed=packed, H=H)
    D = H.shape[0]
    assert packed.dtype == H.dtype == torch.float32 and packed.numel() == lib.gq_upper_packed_floats(D)
    with torch.cuda.device(dev):
        check(lib.gq_unpack_upper(ptr(packed), D, ptr(H), stream_of(H)), 'gq_unpack_upper')


def prep_weight_t(W: torch.Tensor, perm) -> torch.Tensor:
    """Wt (C, R) fp32 with Wt[c, r] = W[r, perm[c]]."""
    dev = require_cuda(W, perm)
    require_contig(W=W, perm=perm)
    R, C_ = W.shape
    Wt = torch.empty(C_, R, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_prep_weight_t(ptr(W), _cabi.DTYPE_CODE[W.dtype], ptr(perm), ptr(Wt), R, C_, stream_of(W)),
              'gq_prep_weight_t')
    return Wt


def prep_weight_t2(W: torch.Tensor, perm, perm_inv=None, want_colsq: bool = True):
    """prep_weight_t plus what the result metrics need later: orig_t (C, R) in W's dtype (the same elements, kept because
    the solver consumes Wt) and colsq (C) fp32 = sum_r W[r, perm[c]]^2.  Returns Wt, orig_t, colsq.
    With perm_inv (HessianHook.perm_inv, or no permutation at all) the scatter-form kernel runs: W read in its own
    order, 128-byte writes (gq_prep_weight_t3); with perm alone, the gather form (gq_prep_weight_t2)."""
    dev = require_cuda(W, perm)
    require_contig(W=W, perm=perm)
    R, C_ = W.shape
    Wt = torch.empty(C_, R, dtype=torch.float32, device=dev)
    orig_t = torch.empty(C_, R, dtype=W.dtype, device=dev)
    colsq = torch.empty(C_, dtype=torch.float32, device=dev) if want_colsq else None
    scatter = (perm is None or perm_inv is not None) and C_ % 2 == 0
    with torch.cuda.device(dev):
        if scatter:
            if perm_inv is not None:
                require_cuda(W, perm_inv)
                perm_inv = perm_inv.to(torch.int64).contiguous()
                assert perm_inv.numel() == C_
            check(lib.gq_prep_weight_t3(ptr(W), _cabi.DTYPE_CODE[W.dtype], ptr(perm_inv) if perm is not None else None,
                                        ptr(Wt), ptr(orig_t), ptr(colsq), R, C_, stream_of(W)), 'gq_prep_weight_t3')
        else:
            check(lib.gq_prep_weight_t2(ptr(W), _cabi.DTYPE_CODE[W.dtype], ptr(perm), ptr(Wt), ptr(orig_t), ptr(colsq), R, C_,
                                        stream_of(W)), 'gq_prep_weight_t2')
    return Wt, orig_t, colsq


def solve_metrics(err_t: torch.Tensor, quant_t: torch.Tensor, orig_t: torch.Tensor) -> torch.Tensor:
    """-> float64 tensor [sum err_t^2, sum (quant_t - orig_t)^2] on the device (one pass over the three arrays)."""
    dev = require_cuda(err_t, quant_t, orig_t)
    require_contig(err_t=err_t, quant_t=quant_t, orig_t=orig_t)
    assert err_t.shape == quant_t.shape == orig_t.shape and err_t.dtype == quant_t.dtype == torch.float32
    out = torch.empty(2, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_solve_metrics(ptr(err_t), ptr(quant_t), ptr(orig_t), _cabi.DTYPE_CODE[orig_t.dtype], err_t.numel(),
                                   ptr(out), stream_of(err_t)), 'gq_solve_metrics')
    return out


def finalize_quant(err_t, quant_t: torch.Tensor, orig_t: torch.Tensor, scale_gr: torch.Tensor, qzero_gr: torch.Tensor, perm, U,
                   row_segments=None):
    """Result assembly behind the solver in one library call (gq_finalize_quant): -> (stats, scale, qzero) with stats a
    float64 device tensor (n_segments, 4) = per row segment [sum err^2, sum (quant - W)^2, sum_j sum_r W_t[j, r]^2 /
    U[perm_j, perm_j]^2, sum dtype(w^2)], scale (R, G) in orig_t.dtype and qzero (R, G) uint8 (the transposes of the solver's
    (G, R) fp32 tables).  row_segments: row counts of the linears that were solved as one stacked weight (default: one)."""
    dev = require_cuda(err_t, quant_t, orig_t, scale_gr, qzero_gr, perm, U)
    require_contig(err_t=err_t, quant_t=quant_t, orig_t=orig_t, scale_gr=scale_gr, qzero_gr=qzero_gr, perm=perm, U=U)
    C_, R = quant_t.shape
    G = scale_gr.shape[0]
    assert orig_t.shape == quant_t.shape and quant_t.dtype == torch.float32
    assert tuple(scale_gr.shape) == (G, R) == tuple(qzero_gr.shape) and scale_gr.dtype == qzero_gr.dtype == torch.float32
    if perm is not None and perm.dtype != torch.int64:
        raise GptqB200Error('perm must be int64')
    rows = [int(R)] if row_segments is None else [int(n) for n in row_segments]
    if sum(rows) != R or min(rows) <= 0:
        raise GptqB200Error('row_segments must add up to the number of rows')
    bounds = [0]
    for n in rows:
        bounds.append(bounds[-1] + n)
    c_bounds = (ctypes.c_int * len(bounds))(*bounds)
    stats = torch.empty(len(rows), 4, dtype=torch.float64, device=dev)
    scale = torch.empty(R, G, dtype=orig_t.dtype, device=dev)
    qzero = torch.empty(R, G, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_finalize_quant(ptr(err_t), ptr(quant_t), ptr(orig_t), _cabi.DTYPE_CODE[orig_t.dtype], R, C_, ptr(scale_gr),
                                    ptr(qzero_gr), G, ptr(scale), ptr(qzero), ptr(perm), ptr(U), c_bounds, len(rows),
                                    ptr(stats), stream_of(quant_t)), 'gq_finalize_quant')
    return stats, scale, qzero


def gptq_solve(Wt: torch.Tensor, U: torch.Tensor, group: int, maxq: float, want_quant: bool = True):
    """K3+K4 in the fp32 semantics of gptq_quant(gptq_use_kernel=False).  Wt (C, R) is
    consumed (updated in place).  Returns quant_t, codes_t (uint8), err_t, scale_gr, qzero_gr."""
    dev = require_cuda(Wt, U)
    require_contig(Wt=Wt, U=U)
    C_, R = Wt.shape
    quant_t = torch.empty_like(Wt) if want_quant else None
    err_t = torch.empty_like(Wt)
    codes_t = torch.empty(C_, R, dtype=torch.uint8, device=dev)
    scale_gr = torch.empty(C_ // group, R, dtype=torch.float32, device=dev)
    qzero_gr = torch.empty(C_ // group, R, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_gptq_solve(ptr(Wt), ptr(U), ptr(quant_t), ptr(codes_t), ptr(err_t), ptr(scale_gr),
                                ptr(qzero_gr), R, C_, int(group), float(maxq), stream_of(Wt)), 'gq_gptq_solve')
    return quant_t, codes_t, err_t, scale_gr, qzero_gr


def gptq_solve_static(Wt: torch.Tensor, U: torch.Tensor, scale_gr: torch.Tensor, qzero_gr: torch.Tensor, group: int, maxq: float):
    """K3+K4 in the kernel-path semantics of gptq_quant(gptq_use_kernel=True): static per-group scale_gr / qzero_gr
    (C/group, R) fp32, fp16-rounded dequant -- the reference's per-group gptq_quantize_range loop in one call.
    Wt (C, R) is consumed.  Returns quant_t, codes_t (int16), err_t."""
    dev = require_cuda(Wt, U, scale_gr, qzero_gr)
    require_contig(Wt=Wt, U=U, scale_gr=scale_gr, qzero_gr=qzero_gr)
    C_, R = Wt.shape
    assert scale_gr.shape == (C_ // group, R) and qzero_gr.shape == (C_ // group, R)
    assert scale_gr.dtype == torch.float32 and qzero_gr.dtype == torch.float32
    quant_t = torch.empty_like(Wt)
    err_t = torch.empty_like(Wt)
    codes_t = torch.empty(C_, R, dtype=torch.int16, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_gptq_solve_static(ptr(Wt), ptr(U), ptr(quant_t), ptr(codes_t), ptr(err_t), ptr(scale_gr),
                                       ptr(qzero_gr), R, C_, int(group), float(maxq), stream_of(Wt)), 'gq_gptq_solve_static')
    return quant_t, codes_t, err_t


def gptq_solve_ex(Wt: torch.Tensor, U: torch.Tensor, scale_gr: torch.Tensor, qzero_gr: torch.Tensor, group: int, maxq_groups,
                  mode: int, a: int = 0, b: int = None, out=None):
    """K3+K4 over columns [a, b) with per-group maxq (list of C/group floats) -- gq_gptq_solve_ex.
    mode 0: dynamic asymmetric parameters (scale_gr / qzero_gr written); 3: static, 16-bit dequant, int16 codes;
    4: static, fp32 arithmetic (symmetric / MSE / EXL2 grids on the fp32 path).  Wt (C, R) is consumed.
    `out` = (quant_t, codes_t, err_t) from an earlier call over another column range, or None to allocate.
    Returns quant_t, codes_t, err_t."""
    import ctypes
    dev = require_cuda(Wt, U, scale_gr, qzero_gr)
    require_contig(Wt=Wt, U=U, scale_gr=scale_gr, qzero_gr=qzero_gr)
    C_, R = Wt.shape
    G = C_ // group
    assert scale_gr.shape == (G, R) and qzero_gr.shape == (G, R)
    assert scale_gr.dtype == torch.float32 and qzero_gr.dtype == torch.float32
    assert len(maxq_groups) == G
    if b is None:
        b = C_
    if out is None:
        out = (torch.empty_like(Wt), torch.empty(C_, R, dtype=torch.int16 if mode == 3 else torch.uint8, device=dev),
               torch.empty_like(Wt))
    quant_t, codes_t, err_t = out
    mq = (ctypes.c_float * G)(*[float(v) for v in maxq_groups])
    with torch.cuda.device(dev):
        check(lib.gq_gptq_solve_ex(ptr(Wt), ptr(U), ptr(quant_t), ptr(codes_t), ptr(err_t), ptr(scale_gr), ptr(qzero_gr),
                                   R, C_, int(group), ctypes.cast(mq, ctypes.c_void_p), 0.0, int(mode), int(a), int(b),
                                   stream_of(Wt)), 'gq_gptq_solve_ex')
    return quant_t, codes_t, err_t


def find_params(Wt: torch.Tensor, group: int, maxq_dev: torch.Tensor, sym: bool, scale_gr: torch.Tensor, qzero_gr: torch.Tensor,
                g0: int = 0, n_groups: int = None) -> None:
    """Quantizer.find_params (quant.py:89-118) for groups [g0, g0 + n_groups) of the transposed weights Wt (C, R):
    writes rows g0.. of scale_gr / qzero_gr (C/group, R) fp32.  maxq_dev: (C/group) fp32 on the device."""
    dev = require_cuda(Wt, maxq_dev, scale_gr, qzero_gr)
    require_contig(Wt=Wt, maxq_dev=maxq_dev, scale_gr=scale_gr, qzero_gr=qzero_gr)
    C_, R = Wt.shape
    G = C_ // group
    if n_groups is None:
        n_groups = G - g0
    assert 0 <= g0 and g0 + n_groups <= G and scale_gr.shape == (G, R) and qzero_gr.shape == (G, R) and maxq_dev.shape == (G,)
    assert Wt.dtype == torch.float32 and maxq_dev.dtype == torch.float32 and scale_gr.dtype == torch.float32
    with torch.cuda.device(dev):
        check(lib.gq_find_params(ptr(Wt[g0 * group:]), R, int(group), int(n_groups), ptr(maxq_dev[g0:]), int(bool(sym)),
                                 ptr(scale_gr[g0:]), ptr(qzero_gr[g0:]), stream_of(Wt)), 'gq_find_params')


def mse_refine(Wt: torch.Tensor, group: int, maxq_dev: torch.Tensor, grid: torch.Tensor, norm: float, scale_gr: torch.Tensor,
               qzero_gr: torch.Tensor, g0: int = 0, n_groups: int = None, want_index: bool = False):
    """Quantizer.mse (quant.py:152-156): scale_gr[g0:g0+n_groups] *= grid[argmin error] per row, in place.
    grid: (Q) fp32 shrink factors on the device.  Returns the chosen indices (n_groups, R) int32 when asked."""
    dev = require_cuda(Wt, maxq_dev, grid, scale_gr, qzero_gr)
    require_contig(Wt=Wt, maxq_dev=maxq_dev, grid=grid, scale_gr=scale_gr, qzero_gr=qzero_gr)
    C_, R = Wt.shape
    G = C_ // group
    if n_groups is None:
        n_groups = G - g0
    assert 0 <= g0 and g0 + n_groups <= G and scale_gr.shape == (G, R) and qzero_gr.shape == (G, R) and maxq_dev.shape == (G,)
    assert grid.dtype == torch.float32 and grid.dim() == 1
    idx = torch.empty(n_groups, R, dtype=torch.int32, device=dev) if want_index else None
    with torch.cuda.device(dev):
        check(lib.gq_mse_refine(ptr(Wt[g0 * group:]), R, int(group), int(n_groups), ptr(maxq_dev[g0:]), ptr(grid), int(grid.numel()),
                                float(norm), ptr(scale_gr[g0:]), ptr(qzero_gr[g0:]), ptr(idx) if idx is not None else None,
                                stream_of(Wt)), 'gq_mse_refine')
    return idx


def _group_args(Wt, maxq_dev, scale_gr, qzero_gr, c0, ncols, n_groups, g0):
    dev = require_cuda(Wt, maxq_dev, scale_gr, qzero_gr)
    require_contig(Wt=Wt, maxq_dev=maxq_dev, scale_gr=scale_gr, qzero_gr=qzero_gr)
    C_, R = Wt.shape
    G = scale_gr.shape[0]
    if not (0 <= c0 and ncols > 0 and n_groups > 0 and c0 + ncols * n_groups <= C_ and 0 <= g0 and g0 + n_groups <= G):
        raise GptqB200Error('column / group range out of bounds')
    if scale_gr.shape != (G, R) or qzero_gr.shape != (G, R) or maxq_dev.shape != (G,):
        raise GptqB200Error('scale_gr / qzero_gr must be (G, R) and maxq_dev (G)')
    if not (Wt.dtype == maxq_dev.dtype == scale_gr.dtype == qzero_gr.dtype == torch.float32):
        raise GptqB200Error('float32 tensors expected')
    return dev, R


def find_params_cols(Wt: torch.Tensor, c0: int, ncols: int, n_groups: int, maxq_dev: torch.Tensor, g0: int, sym: bool,
                     scale_gr: torch.Tensor, qzero_gr: torch.Tensor) -> None:
    """Quantizer.find_params (quant.py:89-118) for n_groups consecutive groups of `ncols` columns each starting at column
    c0 of Wt (C, R); the parameters go to rows g0.. of scale_gr / qzero_gr (G, R) and use maxq_dev[g0..].  Any group size
    (the groups of a linear need not be uniform: one call per run of equal sizes)."""
    dev, R = _group_args(Wt, maxq_dev, scale_gr, qzero_gr, c0, ncols, n_groups, g0)
    with torch.cuda.device(dev):
        check(lib.gq_find_params(ptr(Wt[c0:]), R, int(ncols), int(n_groups), ptr(maxq_dev[g0:]), int(bool(sym)),
                                 ptr(scale_gr[g0:]), ptr(qzero_gr[g0:]), stream_of(Wt)), 'gq_find_params')


def mse_refine_cols(Wt: torch.Tensor, c0: int, ncols: int, n_groups: int, maxq_dev: torch.Tensor, g0: int,
                    grid: torch.Tensor, norm: float, scale_gr: torch.Tensor, qzero_gr: torch.Tensor) -> None:
    """Quantizer.mse (quant.py:152-156) with the error summed over columns [c0 + i * ncols, c0 + (i + 1) * ncols) for
    group g0 + i: the group's own columns, or -- gptq_py.py:178-184 -- the columns of the GPTQ block the group starts in
    (n_groups = 1 then).  ncols <= 1024."""
    dev, R = _group_args(Wt, maxq_dev, scale_gr, qzero_gr, c0, ncols, n_groups, g0)
    require_cuda(Wt, grid)
    require_contig(grid=grid)
    if grid.dtype != torch.float32 or grid.dim() != 1:
        raise GptqB200Error('grid must be a 1-d float32 tensor')
    with torch.cuda.device(dev):
        check(lib.gq_mse_refine(ptr(Wt[c0:]), R, int(ncols), int(n_groups), ptr(maxq_dev[g0:]), ptr(grid), int(grid.numel()),
                                float(norm), ptr(scale_gr[g0:]), ptr(qzero_gr[g0:]), None, stream_of(Wt)), 'gq_mse_refine')


def rtn_quantize(Wt: torch.Tensor, c0: int, ncols: int, n_groups: int, maxq_dev: torch.Tensor, g0: int,
                 scale_gr: torch.Tensor, qzero_gr: torch.Tensor, quant_t: torch.Tensor, codes_t: torch.Tensor) -> None:
    """Round to nearest with the given parameters (gptq_py.py:120-123) for n_groups groups of `ncols` columns from
    column c0: codes_t (C, R) uint8 and quant_t (C, R) fp32 are written for those columns."""
    dev, R = _group_args(Wt, maxq_dev, scale_gr, qzero_gr, c0, ncols, n_groups, g0)
    require_cuda(Wt, quant_t, codes_t)
    require_contig(quant_t=quant_t, codes_t=codes_t)
    if quant_t.shape != Wt.shape or codes_t.shape != Wt.shape or quant_t.dtype != torch.float32 or codes_t.dtype != torch.uint8:
        raise GptqB200Error('quant_t (C, R) float32 and codes_t (C, R) uint8 expected')
    with torch.cuda.device(dev):
        check(lib.gq_rtn_quantize(ptr(Wt[c0:]), R, int(ncols), int(n_groups), ptr(maxq_dev[g0:]), ptr(scale_gr[g0:]),
                                  ptr(qzero_gr[g0:]), ptr(quant_t[c0:]), ptr(codes_t[c0:]), stream_of(Wt)), 'gq_rtn_quantize')


def pack_int4(codes: torch.Tensor) -> torch.Tensor:
    """K5.  codes (R, C) uint8 in [0, 15] -> (R, C/8) int32 words (bit pattern of uint32)."""
    dev = require_cuda(codes)
    require_contig(codes=codes)
    R, C_ = codes.shape
    out = torch.empty(R, C_ // 8, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_pack_int4(ptr(codes), ptr(out), R, C_, stream_of(codes)), 'gq_pack_int4')
    return out


def unpack_int4(packed: torch.Tensor) -> torch.Tensor:
    dev = require_cuda(packed)
    require_contig(packed=packed)
    R, K8 = packed.shape
    out = torch.empty(R, K8 * 8, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_unpack_int4(ptr(packed), ptr(out), R, K8 * 8, stream_of(packed)), 'gq_unpack_int4')
    return out


def transpose_pack_int4(codes_t: torch.Tensor, want_codes: bool = True, want_packed: bool = True):
    """codes_t (C, R) uint8 -> codes (R, C) uint8 and/or packed (R, C/8) int32."""
    dev = require_cuda(codes_t)
    require_contig(codes_t=codes_t)
    C_, R = codes_t.shape
    codes = torch.empty(R, C_, dtype=torch.uint8, device=dev) if want_codes else None
    packed = torch.empty(R, C_ // 8, dtype=torch.int32, device=dev) if want_packed else None
    with torch.cuda.device(dev):
        check(lib.gq_transpose_pack_int4(ptr(codes_t), ptr(codes), ptr(packed), R, C_, stream_of(codes_t)),
              'gq_transpose_pack_int4')
    return codes, packed


def w4a16_gemm(x: torch.Tensor, qweight: torch.Tensor, scale_t: torch.Tensor, qzero_t: torch.Tensor, perm,
               group: int = 128) -> torch.Tensor:
    """K6.  y (M, N) = x[:, perm] @ dequant(qweight)^T;  qweight (N, K/8) int32; scale_t (G, N) in x.dtype and
    qzero_t (G, N) uint8 are group-major (transposes of quant_meta's scale / qzero); perm (K) int32 or None."""
    dev = require_cuda(x, qweight, scale_t, qzero_t, perm)
    require_contig(x=x, qweight=qweight, scale_t=scale_t, qzero_t=qzero_t, perm=perm)
    M, K = x.shape
    N = qweight.shape[0]
    G = K // int(group)
    if (qweight.shape[1] * 8 != K or scale_t.dtype != x.dtype or qzero_t.dtype != torch.uint8
            or tuple(scale_t.shape) != (G, N) or tuple(qzero_t.shape) != (G, N)):
        raise GptqB200Error('w4a16_gemm: inconsistent operands (scale_t / qzero_t must be (K/group, N))')
    if perm is not None and perm.dtype != torch.int32:
        raise GptqB200Error('perm must be int32')
    y = torch.empty(M, N, dtype=x.dtype, device=dev)
    work = torch.empty(lib.gq_w4a16_workspace_bytes(M, N, K), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_w4a16_gemm(ptr(x), ptr(qweight), ptr(scale_t), ptr(qzero_t), ptr(perm), ptr(y), ptr(work), M, N, K,
                                int(group), _cabi.DTYPE_CODE[x.dtype], stream_of(x)), 'gq_w4a16_gemm')
    return y


def w4a16_gemv_static(x: torch.Tensor, qweight: torch.Tensor, scale_t: torch.Tensor, qzero_t: torch.Tensor, perm,
                      out: torch.Tensor, group: int = 128) -> torch.Tensor:
    """Allocation-free batch 1-2 variant of w4a16_gemm (gq_w4a16_gemv_static): the caller passes `out` (M, N); the scratch
    is a fixed 64 KB device array inside the library (M <= 2, K <= 16384; calls that overlap on one GPU race on it)."""
    dev = require_cuda(x, qweight, scale_t, qzero_t, perm, out)
    require_contig(x=x, qweight=qweight, scale_t=scale_t, qzero_t=qzero_t, perm=perm, out=out)
    M, K = x.shape
    N = qweight.shape[0]
    G = K // int(group)
    if (qweight.shape[1] * 8 != K or scale_t.dtype != x.dtype or qzero_t.dtype != torch.uint8 or out.dtype != x.dtype
            or tuple(scale_t.shape) != (G, N) or tuple(qzero_t.shape) != (G, N) or tuple(out.shape) != (M, N)):
        raise GptqB200Error('w4a16_gemv_static: inconsistent operands (scale_t / qzero_t (K/group, N), out (M, N))')
    if perm is not None and perm.dtype != torch.int32:
        raise GptqB200Error('perm must be int32')
    with torch.cuda.device(dev):
        check(lib.gq_w4a16_gemv_static(ptr(x), ptr(qweight), ptr(scale_t), ptr(qzero_t), ptr(perm), ptr(out), M, N, K,
                                       int(group), _cabi.DTYPE_CODE[x.dtype], stream_of(x)), 'gq_w4a16_gemv_static')
    return out


def w4a16_dequant(qweight: torch.Tensor, scale_t: torch.Tensor, qzero_t: torch.Tensor, group: int = 128) -> torch.Tensor:
    """Dense (N, K) reconstruction in scale_t.dtype from the packed format, permuted column order kept:
    dense[n, k] = ((q - z) * s) with the reference's rounding (quant.py:222-237)."""
    dev = require_cuda(qweight, scale_t, qzero_t)
    require_contig(qweight=qweight, scale_t=scale_t, qzero_t=qzero_t)
    N, K = qweight.shape[0], qweight.shape[1] * 8
    G = K // int(group)
    if qzero_t.dtype != torch.uint8 or tuple(scale_t.shape) != (G, N) or tuple(qzero_t.shape) != (G, N):
        raise GptqB200Error('w4a16_dequant: scale_t / qzero_t must be (K/group, N)')
    dense = torch.empty(N, K, dtype=scale_t.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_w4a16_dequant(ptr(qweight), ptr(scale_t), ptr(qzero_t), ptr(dense), N, K, int(group),
                                   _cabi.DTYPE_CODE[scale_t.dtype], stream_of(qweight)), 'gq_w4a16_dequant')
    return dense


def reconstruct_dense(qweight: torch.Tensor, scale: torch.Tensor, qzero: torch.Tensor, perm_inv: torch.Tensor | None,
                      group: int) -> torch.Tensor:
    """Dense (R, C) weight in scale.dtype straight from quant_meta's tensors (one code per byte, (R, G) scale / qzero, int16
    perm_inv or None), uniform group size `group`: construct_matrix + the perm_inv gather of reconstruct_nn_linear
    (quant.py:222-237, 262-279) in one launch, no host synchronisation."""
    dev = require_cuda(qweight, scale, qzero)
    require_contig(qweight=qweight, scale=scale, qzero=qzero)
    R, C = qweight.shape
    G = C // int(group)
    if qweight.dtype != torch.uint8 or qzero.dtype != torch.uint8 or tuple(scale.shape) != (R, G) or tuple(qzero.shape) != (R, G):
        raise GptqB200Error('reconstruct_dense: qweight (R, C) uint8, scale / qzero (R, C / group)')
    if perm_inv is not None:
        if perm_inv.dtype != torch.int16 or perm_inv.numel() != C or not perm_inv.is_contiguous() or perm_inv.device != dev:
            raise GptqB200Error('reconstruct_dense: perm_inv must be a contiguous int16 tensor of C entries on the same device')
    dense = torch.empty(R, C, dtype=scale.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_reconstruct_dense(ptr(qweight), ptr(scale), ptr(qzero), ptr(perm_inv), ptr(dense), R, C, int(group),
                                       _cabi.DTYPE_CODE[scale.dtype], stream_of(qweight)), 'gq_reconstruct_dense')
    return dense


def llama_rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """weight * (x.float() * rsqrt(mean(x.float()^2, -1) + eps)).to(x.dtype) over the last dimension (HF LlamaRMSNorm), one
    pass; x fp16 / bf16 contiguous with at most 8192 entries per row; `out` (same shape, contiguous; may be x) or a new tensor."""
    dev = require_cuda(x, weight)
    require_contig(x=x, weight=weight)
    D = x.shape[-1]
    if weight.dtype != x.dtype or weight.numel() != D:
        raise GptqB200Error('llama_rms_norm: weight must be (D) of x.dtype')
    if out is None:
        out = torch.empty_like(x)
    elif out.shape != x.shape or out.dtype != x.dtype or not out.is_contiguous() or out.device != dev:
        raise GptqB200Error('llama_rms_norm: out must be a contiguous tensor like x')
    with torch.cuda.device(dev):
        check(lib.gq_llama_rms_norm(ptr(x), ptr(weight), ptr(out), x.numel() // max(D, 1), D, float(eps), _cabi.DTYPE_CODE[x.dtype],
                                    stream_of(x)), 'gq_llama_rms_norm')
    return out


def llama_rope_(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, heads: int) -> torch.Tensor:
    """In place: x (..., seqlen, heads * head_dim) <- x * cos + rotate_half(x) * sin per head (HF convention), every operation
    rounded to x.dtype like the PyTorch expression; cos / sin (seqlen, head_dim) of x.dtype."""
    dev = require_cuda(x, cos, sin)
    require_contig(x=x, cos=cos, sin=sin)
    seqlen, hd = cos.shape
    if x.shape[-1] != heads * hd or x.shape[-2] != seqlen or cos.dtype != x.dtype or sin.dtype != x.dtype or sin.shape != cos.shape:
        raise GptqB200Error('llama_rope_: x (..., seqlen, heads * head_dim), cos / sin (seqlen, head_dim) of x.dtype')
    with torch.cuda.device(dev):
        check(lib.gq_llama_rope(ptr(x), ptr(cos), ptr(sin), x.numel() // (heads * hd), int(heads), int(hd), int(seqlen),
                                _cabi.DTYPE_CODE[x.dtype], stream_of(x)), 'gq_llama_rope')
    return x


def llama_silu_mul(gate: torch.Tensor, up: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """F.silu(gate) * up with PyTorch's roundings (silu in fp32 rounded once, the product in the model dtype), one pass."""
    dev = require_cuda(gate, up)
    require_contig(gate=gate, up=up)
    if gate.shape != up.shape or gate.dtype != up.dtype:
        raise GptqB200Error('llama_silu_mul: gate and up must have the same shape and dtype')
    if out is None:
        out = torch.empty_like(gate)
    elif out.shape != gate.shape or out.dtype != gate.dtype or not out.is_contiguous() or out.device != dev:
        raise GptqB200Error('llama_silu_mul: out must be a contiguous tensor like gate')
    with torch.cuda.device(dev):
        check(lib.gq_llama_silu_mul(ptr(gate), ptr(up), ptr(out), gate.numel(), _cabi.DTYPE_CODE[gate.dtype], stream_of(gate)),
              'gq_llama_silu_mul')
    return out


def gemm_f16_nt(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """y (M, N) = x (M, K) @ w (N, K)^T in x.dtype (fp16 / bf16), fp32 accumulation on tcgen05 CTA pairs: the dense GEMM of
    K6's large-batch route (the reference's nn.Linear on the reconstructed weights, llama.py:172-177)."""
    dev = require_cuda(x, w)
    require_contig(x=x, w=w)
    if x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1] or x.dtype != w.dtype or x.dtype not in (torch.float16, torch.bfloat16):
        raise GptqB200Error(f'gemm_f16_nt: x {tuple(x.shape)} {x.dtype} / w {tuple(w.shape)} {w.dtype}')
    M, K = x.shape
    N = w.shape[0]
    y = torch.empty(M, N, dtype=x.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib.gq_gemm_f16_nt(ptr(x), ptr(w), ptr(y), M, N, K, _cabi.DTYPE_CODE[x.dtype], stream_of(x)), 'gq_gemm_f16_nt')
    return y


def sgemm(A, B, C_, M, N, K, lda, ldb, ldc, alpha, beta, flags) -> None:
    """Test hook for the fp32 GEMM building block (operands are views into larger buffers)."""
    dev = require_cuda(A, B, C_)
    with torch.cuda.device(dev):
        check(lib.gq_sgemm(ptr(A), ptr(B), ptr(C_), M, N, K, lda, ldb, ldc, float(alpha), float(beta), int(flags),
                           stream_of(C_)), 'gq_sgemm')


def gemm_tf32x3(A, B, C_, M, N, K, lda, ldb, ldc, alpha, upper_only=0) -> None:
    """C (M, N) += alpha * A^T B on the tensor cores with a 3 x TF32 split (A stored (K, M), B (K, N))."""
    dev = require_cuda(A, B, C_)
    with torch.cuda.device(dev):
        check(lib.gq_gemm_tf32x3(ptr(A), ptr(B), ptr(C_), M, N, K, lda, ldb, ldc, float(alpha), int(upper_only),
                                 stream_of(C_)), 'gq_gemm_tf32x3')


def set_pdl(on: int) -> int:
    """Programmatic dependent launch on the K2 / K3+K4 kernel chains (1 = default); returns the previous value."""
    return int(lib.gq_set_pdl(int(on)))


def set_w4a16_path(path: int) -> int:
    """K6 dispatch hook (0 auto, 1 never the mma.sync kernels, 2 the experimental weights-on-N tcgen05 kernel for batch <= 128,
    3 auto without the batch 9..16 kernel, 4 that kernel wherever it supports the shape); returns the old value."""
    return int(lib.gq_set_w4a16_path(int(path)))


def set_syrk_kchunk(kblocks: int) -> int:
    """K1: 64-token k-blocks accumulated in tensor memory per partial sum (0 = a whole call); returns the old value."""
    return int(lib.gq_set_syrk_kchunk(int(kblocks)))


def set_tf32_kchunk(kblocks: int) -> int:
    """3xTF32 GEMM: k-blocks (of 16) accumulated in tensor memory per partial sum (0 = unlimited); returns the old value."""
    return int(lib.gq_set_tf32_kchunk(int(kblocks)))


def set_solve_fused(on) -> int:
    """A/B hook: fused super-block solver (default) vs one launch per block + near-update GEMMs; returns the old value."""
    return int(lib.gq_set_solve_fused(int(bool(on))))


def set_solve_super(blocks: int) -> int:
    """Blocks per super-block of K4's two-level update (1 = one full trailing update per block); returns the old value."""
    return int(lib.gq_set_solve_super(int(blocks)))


def set_tf32x3(enable) -> int:
    """A/B hook: 0 = fp32 SIMT GEMM, 1 = 3xTF32 with the round-to-nearest split, 2 = 3xTF32 with the truncation split
    (hi is the landed tile itself), 3 (default) = 2 plus CTA-pair 256 x 256 tiles for multi-wave launches; returns the
    previous setting."""
    return int(lib.gq_set_tf32x3(int(enable)))
