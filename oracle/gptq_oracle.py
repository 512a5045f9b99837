"""CPU oracle for the GPTQ quantization hot path  --  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the arithmetic of the reference's fp32 PyTorch
GPTQ path so that the CUDA kernels can be checked against it on a box where
/root/reference does not exist.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import it.  The product
(gptq-v2_b200/) never does: it fails loudly when the CUDA library is missing.

Why torch-CPU and not numpy: the reference's own CPU path *is* torch
(torch.addmm, torch.linalg.cholesky_ex, torch.cholesky_inverse, Tensor.round),
so using the same primitives keeps the restatement bit-identical to the
reference run on the same host (checked by oracle/gen_golden.py, which imports
the real reference in the build container and freezes outputs under
tests/golden/).

Parity pinning: the reference holds NO golden vectors for this path
(SURVEY.md section 4).  The oracle is pinned instead against outputs of the
reference itself (gptq_py.py + quant.py imported from /root/reference with a
stub `gptq` package) -- see oracle/gen_golden.py and tests/test_oracle_golden.py.
The factorisation goes through LAPACK (un-vendored, torch-bundled): parity for
U = chol(H^-1) is pinned only through those frozen outputs.

Every function cites the reference lines it follows (paths relative to
/root/reference).
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch

EPS = 1e-12  # quant.py:7


# --------------------------------------------------------------------------
# a1/a2  Hessian accumulation                      gptq_py.py:19-36
# --------------------------------------------------------------------------
def hessian_accumulate(H: Optional[torch.Tensor], n_seq: int, x: torch.Tensor):
    """H += X^T X in fp32; n_seq counts *sequences* (gptq_py.py:25-33).

    x: (B, N, D) or (N, D) 16-bit or fp32.  Returns (H, n_seq).
    """
    if x.dim() <= 2:
        x = x[None]
    n_seq += x.shape[0]
    d = x.shape[-1]
    if H is None:
        H = torch.zeros(d, d, dtype=torch.float32)
    x2 = x.reshape(-1, d).to(torch.float32)
    H.addmm_(x2.t(), x2)  # beta=1, alpha=1  (gptq_py.py:33)
    return H, n_seq


# --------------------------------------------------------------------------
# a3  damp + Cholesky + inverse factor             gptq_py.py:39-66
# --------------------------------------------------------------------------
def hessian_invert(H: torch.Tensor, n_seq: int, damp_ratio: float = 1e-2,
                   act_order: bool = True, perm: Optional[torch.Tensor] = None):
    """Returns dict(H=<scaled, permuted, damped H>, U, perm, perm_inv, tries).

    U is upper triangular with U^T U = H^-1.  `perm` may be injected to
    decouple a test from argsort tie order.
    """
    H = H.clone()
    H *= 2.0 / n_seq                                   # :41
    dead = H.diag() == 0.0                             # :43
    H[dead, dead] = 1.0                                # :44
    perm_inv = None
    if act_order:
        if perm is None:
            perm = H.diag().argsort(descending=True)   # :47
        H = H[perm][:, perm]                           # :48
        perm_inv = perm.argsort()                      # :49
    else:
        perm = None
    idx = torch.arange(H.shape[0])
    damp = damp_ratio * H.diag().mean()                # :52
    U = torch.empty_like(H)
    info = torch.empty((), dtype=torch.int32)
    tries = 0
    ok = False
    for _ in range(100):                               # :55-65
        tries += 1
        H[idx, idx] += damp                            # cumulative (:58)
        torch.linalg.cholesky_ex(H, upper=False, check_errors=False, out=(U, info))
        torch.cholesky_inverse(U, upper=False, out=U)
        torch.linalg.cholesky_ex(U, upper=True, check_errors=False, out=(U, info))
        if not U.isnan().any():
            ok = True
            break
    assert ok
    return dict(H=H, U=U, perm=perm, perm_inv=perm_inv, tries=tries)


# --------------------------------------------------------------------------
# a4  scalar quantisation math                     quant.py:10-30, 89-118
# --------------------------------------------------------------------------
def find_params_asym(x: torch.Tensor, bits: int):
    """Per-row asymmetric min/max scale and zero over dim -1 (quant.py:98-118)."""
    maxq = torch.tensor(2.0 ** bits - 1.0, dtype=torch.float32)
    xmax = x.max(dim=-1, keepdim=True)[0].relu()
    xmin = -(-x.min(dim=-1, keepdim=True)[0]).relu()
    scale = (xmax - xmin) / maxq + EPS
    qzero = (-xmin / scale).round()
    return scale, qzero, maxq


def quantize(x, scale, qzero, maxq):
    """quant.py:19"""
    return ((x / scale).round() + qzero).clamp(0.0, maxq)


def dequantize(q, scale, qzero):
    """quant.py:30"""
    return (q - qzero) * scale


# --------------------------------------------------------------------------
# a5  pure-fp32 blocked GPTQ solver (the parity oracle)   gptq_py.py:158-193
# --------------------------------------------------------------------------
def gptq_solve_pure(weight: torch.Tensor, U: torch.Tensor, perm: Optional[torch.Tensor],
                    group_size: int = 128, bits: int = 4, block_size: Optional[int] = None):
    """weight (R, C) any float.  Returns dict(quant, qweight, error, scale, qzero)
    in *permuted* column order, fp32; scale/qzero are (R, G).

    Dynamic group parameters: min/max is taken at each group start on the
    weights that already carry every earlier lazy-batch update
    (gptq_py.py:169-176, 193).
    """
    block_size = block_size or group_size
    W = weight.to(torch.float32).clone()               # :90
    if perm is not None:
        W = W[:, perm]                                 # :93-94
    R, C = W.shape
    assert C % group_size == 0 and C % block_size == 0
    quant = torch.empty_like(W)
    qweight = torch.empty_like(W)
    error = torch.empty_like(W)
    scales, zeros = [], []
    s = z = mq = None
    for i1 in range(0, C, block_size):
        i2 = i1 + block_size
        blk = W[:, i1:i2].clone()                      # :166
        eblk = torch.empty_like(blk)
        for j in range(i1, i2):
            if j % group_size == 0:                    # :169
                s, z, mq = find_params_asym(W[:, j:j + group_size], bits)
                scales.append(s)
                zeros.append(z)
            w = blk[:, j - i1:j - i1 + 1]
            qw = quantize(w, s, z, mq)                 # :188
            q = dequantize(qw, s, z)                   # :189
            err = (w - q) / U[j, j]                    # :190
            qweight[:, j:j + 1] = qw
            quant[:, j:j + 1] = q
            error[:, j:j + 1] = err
            eblk[:, j - i1:j - i1 + 1] = err
            blk[:, j - i1 + 1:] -= err @ U[j:j + 1, j + 1:i2]   # :192
        W[:, i2:] -= eblk @ U[i1:i2, i2:]              # :193
    return dict(quant=quant, qweight=qweight, error=error,
                scale=torch.cat(scales, dim=-1), qzero=torch.cat(zeros, dim=-1))


# --------------------------------------------------------------------------
# a6  kernel path semantics (static scales, fp16 dequant)
#     gptq_py.py:125-157, package/gptq/gptq_cpp.cpp:33-84,
#     package/gptq/gptq_cpp_kernel.cu:21-74 and :113-137
# --------------------------------------------------------------------------
def static_group_params(weight: torch.Tensor, perm: Optional[torch.Tensor],
                        group_size: int = 128, bits: int = 4):
    """Scale/zero from the ORIGINAL (permuted, un-compensated) weights, as the
    reference's kernel path computes them (gptq_py.py:131-134)."""
    W = weight.to(torch.float32)
    if perm is not None:
        W = W[:, perm]
    R, C = W.shape
    s, z, mq = find_params_asym(W.reshape(R, C // group_size, group_size), bits)
    return s[..., 0].contiguous(), z[..., 0].contiguous(), float(mq)


def quantize_range_ref(quant_t, scale, out_q, qzero, maxq, U, weights_t, error_t, a, b):
    """Restatement of quantize_range() for columns [a, b) on the transposed
    (C, R) layout; all tensors mutated in place like the reference.

    Per column c (gptq_cpp_kernel.cu:41-73): q = clamp(rint(w/s)+z, 0, maxq);
    the dequantised value is rounded through fp16 (s, q, z each to half, then
    half subtract, half multiply); err = (w - deq) / U[c,c].
    In-block update (:113-137) is z -= y*x, contracted to one FMA by nvcc.
    Trailing update (gptq_cpp.cpp:81-83) is an fp32 addmm.
    """
    sc = scale.to(torch.float32)
    zp = qzero.to(torch.float32)
    s16 = sc.to(torch.float16)
    z16 = zp.to(torch.float16)
    for c in range(a, b):
        w = weights_t[c]
        x = ((w / sc).round() + zp).clamp(0.0, maxq)
        if out_q is not None:
            out_q[c] = x.to(out_q.dtype)
        h = (x.to(torch.float16) - z16)                # __hsub
        h = h * s16                                    # __hmul
        q = h.to(torch.float32)
        quant_t[c] = q
        e = (w - q) / U[c, c]
        error_t[c] = e
        if c + 1 < b:
            # fma(-e, u, w): emulate the single rounding in float64
            # (24-bit x 24-bit products are exact in float64; the sum of an
            # exact product and a float32 is rounded once when cast back...
            # not always exactly once, so do it properly below).
            u = U[c, c + 1:b]
            weights_t[c + 1:b] = _fma32(-u[:, None], e[None, :], weights_t[c + 1:b])
    if b < weights_t.shape[0]:
        x = U[a:b, b:].t()
        weights_t[b:].addmm_(x, error_t[a:b], beta=1.0, alpha=-1.0)


def _fma32(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """Correctly rounded fp32 fma(a, b, c) on tensors.

    a*b is exact in float64 (48 significant bits); the float64 sum p + c can
    itself round (needs up to ~77 bits), so a plain double evaluation can
    double-round.  Use the round-to-odd trick: compute the float64 sum, detect
    inexactness via TwoSum, force the sticky bit (LSB |= 1) when inexact, then
    round to float32 once.  (Boldo & Melquiond 2008.)
    """
    p = a.to(torch.float64) * b.to(torch.float64)
    c64 = c.to(torch.float64).expand_as(p)
    s = p + c64
    # TwoSum error term
    bb = s - p
    err = (p - (s - bb)) + (c64 - bb)
    si = s.contiguous().view(torch.int64)
    inexact = err != 0
    # round-to-odd: if inexact and LSB even, move one ulp toward the error sign
    lsb_even = (si & 1) == 0
    need = inexact & lsb_even
    # direction: if err has same sign as s -> magnitude increases -> +1 ulp,
    # else magnitude decreases -> the truncated value is one ulp below: -1 ulp
    same = (err > 0) == (s > 0)
    adj = torch.where(same, torch.ones_like(si), -torch.ones_like(si))
    si = torch.where(need, si + adj, si)
    s_odd = si.view(torch.float64)
    return s_odd.to(torch.float32)


def gptq_solve_kernelpath(weight: torch.Tensor, U: torch.Tensor, perm: Optional[torch.Tensor],
                          group_size: int = 128, bits: int = 4):
    """End-to-end restatement of the reference's `gptq_use_kernel=True` branch
    (gptq_py.py:125-157): transposed clone, static per-group parameters, one
    quantize_range per group.  Same return convention as gptq_solve_pure."""
    W = weight.to(torch.float32).clone()
    if perm is not None:
        W = W[:, perm]
    R, C = W.shape
    scale, qzero, maxq = static_group_params(weight, perm, group_size, bits)
    wt = W.t().contiguous()                            # :128
    quant_t = torch.empty_like(wt)
    err_t = torch.empty_like(wt)
    q_t = torch.empty_like(wt, dtype=torch.int16)
    for g, i1 in enumerate(range(0, C, group_size)):
        quantize_range_ref(quant_t, scale[:, g].contiguous(), q_t, qzero[:, g].contiguous(), maxq,
                           U, wt, err_t, i1, i1 + group_size)
    return dict(quant=quant_t.t(), qweight=q_t.t().to(torch.float32), error=err_t.t(),
                scale=scale, qzero=qzero)


# --------------------------------------------------------------------------
# a7  metrics + storage dtypes                     gptq_py.py:195-251
# --------------------------------------------------------------------------
def assemble_result(weight_ref: torch.Tensor, solved: dict, U: torch.Tensor,
                    perm: Optional[torch.Tensor], group_size: int, bits: int):
    """Builds {'quant_meta', 'metrics'} with the reference's keys, dtypes and
    column order (codes stay in permuted order; scale is cast to the weight
    dtype; perm_inv is int16)."""
    R, C = weight_ref.shape
    G = C // group_size
    dtype = weight_ref.dtype
    err = solved['error']
    metrics = {
        'gptq_error': (err ** 2.0).mean().item(),                                  # :196
        # NB mixes un-permuted W with permuted diag(U) -- reproduced as is (:198)
        'gptq_norm': ((weight_ref / U.diag()) ** 2.0).mean().item(),
    }
    if perm is not None:
        p16 = perm.to(torch.int16)
    else:
        p16 = torch.arange(C, dtype=torch.int16)
    perm_inv = p16.argsort().to(torch.int16)                                        # :226
    quant = solved['quant'][:, perm_inv.to(torch.int32)]                            # :231
    diff = ((quant.to(torch.float32) - weight_ref.to(torch.float32)).abs() ** 2.0).mean()
    metrics['l2.0_error'] = diff.item()
    metrics['l2.0_norm'] = (weight_ref.abs() ** 2.0).mean().item()
    meta = {
        'qweight': solved['qweight'].to(torch.uint8),
        'scale': solved['scale'].to(dtype),
        'qzero': solved['qzero'].to(torch.uint8),
        'qscale': None, 'sscale': None,
        'perm_inv': perm_inv,
        'group_sizes': torch.full((G,), group_size, dtype=torch.int16),
        'group_bit_widths': torch.full((G,), bits, dtype=torch.uint8),
        'scale_bit_width': None,
    }
    return {'quant_meta': meta, 'metrics': metrics}


# --------------------------------------------------------------------------
# a8  dense reconstruction + forward               quant.py:222-237, 262-279
# --------------------------------------------------------------------------
def reconstruct_dense(meta: dict, dtype: torch.dtype) -> torch.Tensor:
    """W_hat (R, C) in the model dtype, original column order.  The arithmetic
    is int16 subtract then multiply by the 16-bit scale *in that dtype*."""
    q = meta['qweight'].to(torch.int16)
    z = meta['qzero'].to(torch.int16)
    s = meta['scale'].to(dtype)
    gs = [int(v) for v in meta['group_sizes'].tolist()]
    out = torch.empty(q.shape, dtype=dtype)
    i = 0
    for g, n in enumerate(gs):
        out[:, i:i + n] = (q[:, i:i + n] - z[:, g:g + 1]) * s[:, g:g + 1]
        i += n
    if meta.get('perm_inv') is not None:
        out = out[:, meta['perm_inv'].to(torch.int32)]
    return out


def linear_forward_dense(x: torch.Tensor, meta: dict, dtype: torch.dtype) -> torch.Tensor:
    """y = x W_hat^T, evaluated in fp32 on the dtype-rounded W_hat (the CPU
    stand-in for the reference's cuBLAS HGEMM, llama.py:172-177)."""
    w = reconstruct_dense(meta, dtype).to(torch.float32)
    return x.to(torch.float32) @ w.t()


# --------------------------------------------------------------------------
# K5  int4 / int3-in-nibble packing  (format defined by this build)
# --------------------------------------------------------------------------
def pack_int4(codes: np.ndarray) -> np.ndarray:
    """codes (R, C) uint8 in [0, 15], C % 8 == 0  ->  (R, C/8) uint32.
    Column 8*k + i of a row sits in bits [4i, 4i+4) of word k."""
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    R, C = codes.shape
    assert C % 8 == 0
    c = codes.reshape(R, C // 8, 8).astype(np.uint32)
    out = np.zeros((R, C // 8), dtype=np.uint32)
    for i in range(8):
        out |= (c[:, :, i] & 0xF) << np.uint32(4 * i)
    return out


def unpack_int4(words: np.ndarray) -> np.ndarray:
    words = np.ascontiguousarray(words, dtype=np.uint32)
    R, K = words.shape
    out = np.empty((R, K, 8), dtype=np.uint8)
    for i in range(8):
        out[:, :, i] = (words >> np.uint32(4 * i)) & 0xF
    return out.reshape(R, K * 8)


def w4a16_reference(x: torch.Tensor, qpacked: np.ndarray, scale: torch.Tensor, qzero: torch.Tensor,
                    perm: Optional[torch.Tensor], group_size: int = 128) -> torch.Tensor:
    """y (M, R) fp32 = x[:, perm] @ dequant(codes)^T with dequant in fp32 from the
    16-bit scale.  Independent of reconstruct_dense (no fp16 rounding of W_hat);
    agrees with it to ~1e-3 relative."""
    codes = torch.from_numpy(unpack_int4(qpacked).astype(np.float32))
    R, C = codes.shape
    s = scale.to(torch.float32).repeat_interleave(group_size, dim=1)
    z = qzero.to(torch.float32).repeat_interleave(group_size, dim=1)
    w = (codes - z) * s
    xp = x.to(torch.float32)
    if perm is not None:
        xp = xp[:, perm.to(torch.int64)]
    return xp @ w.t()


# --------------------------------------------------------------------------
# parity metrics used by the tests and bench
# --------------------------------------------------------------------------
def rel_fro(a: torch.Tensor, b: torch.Tensor) -> float:
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def recon_error(W: torch.Tensor, quant_perm: torch.Tensor, Hperm: torch.Tensor,
                perm: Optional[torch.Tensor]) -> float:
    """||W X - W_hat X||_F^2 up to a constant = tr(D H D^T), D = W_perm - W_hat_perm."""
    Wp = W.to(torch.float64)
    if perm is not None:
        Wp = Wp[:, perm]
    Dm = Wp - quant_perm.to(torch.float64)
    return (Dm @ Hperm.to(torch.float64) * Dm).sum().item()


def code_match(q_a: torch.Tensor, q_b: torch.Tensor) -> float:
    return (q_a.to(torch.int32) == q_b.to(torch.int32)).float().mean().item()


# --------------------------------------------------------------------------
# synthetic inputs shared by tests, bench and the golden generator
# 
# --------------------------------------------------------------------------
def synth_linear(R: int, C: int, n_seq: int, seqlen: int, dtype=torch.float16, seed: int = 0):
    g0 = torch.Generator().manual_seed(seed)
    g1 = torch.Generator().manual_seed(seed + 1)
    g2 = torch.Generator().manual_seed(seed + 2)
    colscale = torch.exp(torch.rand(C, generator=g2) * 3.0 - 1.5)
    X = (torch.randn(n_seq, seqlen, C, generator=g0) * colscale).to(dtype)
    W = (torch.randn(R, C, generator=g1) * 0.02).to(dtype)
    return W, X


def run_linear(W: torch.Tensor, X: torch.Tensor, group_size=128, bits=4, damp=1e-2,
               act_order=True, mode: str = 'pure', batch: int = 16, timings: Optional[dict] = None):
    """Whole path for one linear on the CPU: accumulate -> invert -> solve -> assemble."""
    import time
    t0 = time.perf_counter()
    H, n = None, 0
    for i in range(0, X.shape[0], batch):
        H, n = hessian_accumulate(H, n, X[i:i + batch])
    t1 = time.perf_counter()
    inv = hessian_invert(H, n, damp, act_order)
    t2 = time.perf_counter()
    fn = gptq_solve_pure if mode == 'pure' else gptq_solve_kernelpath
    sol = fn(W, inv['U'], inv['perm'], group_size, bits)
    res = assemble_result(W, sol, inv['U'], inv['perm'], group_size, bits)
    t3 = time.perf_counter()
    if timings is not None:
        timings.update(accumulate=t1 - t0, invert=t2 - t1, solve=t3 - t2)
    return dict(H_raw=H, n_seq=n, inv=inv, solved=sol, result=res)
