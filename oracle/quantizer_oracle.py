"""CPU oracle for the remaining Quantizer features  --  TEST INFRASTRUCTURE ONLY.

Symmetric grids, the MSE grid search, EXL2 quantised scales (qscale / sscale) and per-group bit widths, restated
from /root/reference/quant.py and the three branches of gptq_quant (/root/reference/gptq_py.py:101-193) that consume
them.  Same rules as oracle/gptq_oracle.py: only tests/, smoke() and bench.py's CPU legs may import this file.

Parity pinning: oracle/gen_golden.py runs the real reference (round-to-nearest and fp32 GPTQ branches, which need no
CUDA extension) on the cases listed there and freezes its outputs as tests/golden/q_*.npz;
tests/test_oracle_golden.py checks this file against them.  The reference's kernel branch with these options cannot run
here; quantize_range_ref (gptq_oracle.py) stays its restatement.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from oracle import gptq_oracle as G

EPS = G.EPS


def quantize2(x, scale, maxq):
    """quant.py:40-53: quadratic grid for a column of scales."""
    return (x / scale).sqrt().round().clamp(1.0, maxq + 1.0)


def dequantize2(qx, scale):
    """quant.py:56-65"""
    return qx * qx * scale


def find_params(x: torch.Tensor, bits: int, sym: bool = False, scale_bits: Optional[int] = None) -> dict:
    """quant.py:89-118.  x (R, C) fp32.  Returns scale (R,1), qzero (R,1), maxq (), and for EXL2 qscale (R,1),
    sscale (1,1), smaxq ()."""
    if scale_bits is not None and scale_bits > 0:
        sym = True                                                          # :99-100
    maxq = torch.tensor(2.0 ** bits - 1.0, dtype=torch.float32)             # :102
    P = {'maxq': maxq, 'qscale': None, 'sscale': None, 'smaxq': None}
    if sym:
        P['scale'] = x.abs().max(dim=-1, keepdim=True)[0] * (2.0 / maxq) + EPS           # :105
        P['qzero'] = torch.full_like(P['scale'], ((maxq + 1.0) * 0.5).round())           # :106
        if scale_bits is not None and scale_bits > 0:
            smaxq = torch.tensor(2.0 ** scale_bits - 1.0, dtype=torch.float32)           # :109
            P['smaxq'] = smaxq
            P['sscale'] = P['scale'].max(dim=-2, keepdim=True)[0] / (smaxq + 1.0) ** 2.0 + EPS   # :110
            P['qscale'] = quantize2(P['scale'], P['sscale'], smaxq)                       # :111
            P['scale'] = dequantize2(P['qscale'], P['sscale'])                            # :112
    else:
        s, z, _ = G.find_params_asym(x, bits)                                             # :114-118
        P['scale'], P['qzero'] = s, z
    return P


def shrink_grid(max_shrink: float, n_grid: int) -> torch.Tensor:
    """quant.py:152: the candidate shrink factors, 1 first."""
    return 1.0 - torch.arange(0.0, max_shrink, 1.0 / n_grid)


def mse_refine(P: dict, x: torch.Tensor, max_shrink: float = 0.8, n_grid: int = 100, norm: float = 2.4) -> dict:
    """quant.py:152-165 (use_kernel=False): per row, the shrink factor with the smallest sum |deq(q(x)) - x|^norm;
    for EXL2 then one factor for the group's sscale."""
    p = shrink_grid(max_shrink, n_grid)
    q = G.dequantize(G.quantize(x[..., None], P['scale'][..., None] * p, P['qzero'][..., None], P['maxq']),
                     P['scale'][..., None] * p, P['qzero'][..., None])                    # (R, C, Q)   :153
    arg = (q - x[..., None]).abs().pow(norm).sum(dim=-2).argmin(dim=-1, keepdim=True)    # (R, 1)      :155
    P['scale'] = P['scale'] * p[arg]                                                      # :156
    P['mse_index'] = arg
    if P['sscale'] is not None:
        smaxq = P['smaxq']
        P['sscale'] = P['scale'].max(dim=-2, keepdim=True)[0] / (smaxq + 1.0) ** 2.0 + EPS          # :158
        q = dequantize2(quantize2(P['scale'][..., None], P['sscale'][..., None] * p, smaxq),
                        P['sscale'][..., None] * p)                                       # (R, 1, Q)   :160
        arg = (q - P['scale'][..., None]).abs().pow(norm).sum(dim=-3).argmin(dim=-1, keepdim=True)  # (1, 1)  :162
        P['sscale'] = P['sscale'] * p[arg]                                                # :164
        P['qscale'] = quantize2(P['scale'], P['sscale'], smaxq)                           # :165
        P['scale'] = dequantize2(P['qscale'], P['sscale'])                                # :166
    return P


def _group_params(x, bits, opts):
    P = find_params(x, bits, opts.get('sym', False), opts.get('scale_bits'))
    if opts.get('mse', False):
        mse_refine(P, x, opts.get('max_shrink', 0.8), opts.get('n_grid', 100), opts.get('norm', 2.4))
    return P


def _collect(params: Sequence[dict]) -> dict:
    """collate_quantizers (quant.py:186-219) for the fields quant_meta keeps."""
    out = {'scale': torch.cat([p['scale'] for p in params], dim=-1),
           'qzero': torch.cat([p['qzero'] for p in params], dim=-1)}
    if params[0]['sscale'] is not None:
        out['qscale'] = torch.cat([p['qscale'] for p in params], dim=-1)
        out['sscale'] = torch.cat([p['sscale'] for p in params], dim=-1)      # (1, G)
    return out


def solve_general(weight: torch.Tensor, U: Optional[torch.Tensor], perm: Optional[torch.Tensor], group_size: int,
                  bits: Sequence[int], branch: str = 'pure', **opts) -> dict:
    """The three branches of gptq_quant with every Quantizer option.
    branch 'rtn' (gptq_py.py:101-124), 'kernel' (:125-157, through quantize_range_ref), 'pure' (:158-193).
    Group g covers columns [g*group_size, (g+1)*group_size) and is quantised to bits[g]; block size = group size.
    Returns quant / qweight / error (R, C) in permuted order plus the collected per-group parameters."""
    W = weight.to(torch.float32).clone()
    if perm is not None:
        W = W[:, perm]
    R, C = W.shape
    n_groups = C // group_size
    assert n_groups * group_size == C and len(bits) == n_groups
    quant = torch.empty_like(W)
    qweight = torch.empty_like(W)
    params = []
    if branch == 'rtn':
        for g in range(n_groups):
            i1, i2 = g * group_size, (g + 1) * group_size
            P = _group_params(W[:, i1:i2], bits[g], opts)
            params.append(P)
            qweight[:, i1:i2] = qw = G.quantize(W[:, i1:i2], P['scale'], P['qzero'], P['maxq'])
            quant[:, i1:i2] = G.dequantize(qw, P['scale'], P['qzero'])
        error = None
    elif branch == 'kernel':
        wt = W.t().contiguous()
        quant_t, err_t = torch.empty_like(wt), torch.empty_like(wt)
        q_t = torch.empty_like(wt, dtype=torch.int16)
        for g in range(n_groups):
            i1, i2 = g * group_size, (g + 1) * group_size
            P = _group_params(W[:, i1:i2], bits[g], opts)          # from the un-compensated weights (:131-134)
            params.append(P)
            G.quantize_range_ref(quant_t, P['scale'][:, 0].contiguous(), q_t, P['qzero'][:, 0].contiguous(),
                                 float(P['maxq']), U, wt, err_t, i1, i2)
        quant, qweight, error = quant_t.t(), q_t.t().to(torch.float32), err_t.t()
    else:
        error = torch.empty_like(W)
        for g in range(n_groups):
            i1, i2 = g * group_size, (g + 1) * group_size
            P = _group_params(W[:, i1:i2], bits[g], opts)          # from the compensated weights (:171-185)
            params.append(P)
            blk = W[:, i1:i2].clone()
            eblk = torch.empty_like(blk)
            for j in range(i1, i2):
                w = blk[:, j - i1:j - i1 + 1]
                qw = G.quantize(w, P['scale'], P['qzero'], P['maxq'])
                q = G.dequantize(qw, P['scale'], P['qzero'])
                err = (w - q) / U[j, j]
                qweight[:, j:j + 1], quant[:, j:j + 1] = qw, q
                error[:, j:j + 1] = eblk[:, j - i1:j - i1 + 1] = err
                blk[:, j - i1 + 1:] -= err @ U[j:j + 1, j + 1:i2]
            W[:, i2:] -= eblk @ U[i1:i2, i2:]
    out = dict(quant=quant, qweight=qweight, error=error, mse_index=None)
    out.update(_collect(params))
    if opts.get('mse', False):
        out['mse_index'] = torch.cat([p['mse_index'] for p in params], dim=-1)
    return out


def solve_blocks(weight: torch.Tensor, U: torch.Tensor, perm: Optional[torch.Tensor], group_sizes: Sequence[int],
                 bits: Sequence[int], block_sizes: Sequence[int], **opts) -> dict:
    """The fp32 branch of gptq_quant (gptq_py.py:158-193) with GPTQ blocks that differ from the quantisation groups and
    group sizes that need not be uniform, exactly as the reference walks it:
    the parameters of a group are computed when the column loop reaches the group's first column, from the GLOBAL
    weight matrix (:175-176) -- which holds the lazy-batch updates of the finished blocks but NOT the in-block updates
    of the current one -- while the columns themselves are quantised from the error-compensated block copy (:186);
    a quantiser stays in force across block ends (:167,171).  opts: sym, scale_bits, mse (+ max_shrink, n_grid, norm);
    the MSE search runs over the columns of the current BLOCK of the global matrix (:178-184), not over the group's."""
    W = weight.to(torch.float32).clone()
    if perm is not None:
        W = W[:, perm]
    R, C = W.shape
    assert sum(group_sizes) == C and sum(block_sizes) == C and len(bits) == len(group_sizes)
    gid = [0]
    for n in group_sizes:
        gid.append(gid[-1] + int(n))
    bid = [0]
    for n in block_sizes:
        bid.append(bid[-1] + int(n))
    quant, qweight, error = torch.empty_like(W), torch.empty_like(W), torch.empty_like(W)
    params = []
    cur, P = 0, None
    for i1, i2 in zip(bid[:-1], bid[1:]):
        blk = W[:, i1:i2].clone()
        eblk = torch.empty_like(blk)
        for j in range(i1, i2):
            if cur < len(group_sizes) and j == gid[cur]:
                P = find_params(W[:, j:j + int(group_sizes[cur])], int(bits[cur]), opts.get('sym', False),
                                opts.get('scale_bits'))                                      # :172-177
                if opts.get('mse', False):
                    mse_refine(P, W[:, i1:i2], opts.get('max_shrink', 0.8), opts.get('n_grid', 100),
                               opts.get('norm', 2.4))                                        # :178-184
                params.append(P)
                cur += 1
            w = blk[:, j - i1:j - i1 + 1]
            qw = G.quantize(w, P['scale'], P['qzero'], P['maxq'])
            q = G.dequantize(qw, P['scale'], P['qzero'])
            err = (w - q) / U[j, j]
            qweight[:, j:j + 1], quant[:, j:j + 1] = qw, q
            error[:, j:j + 1] = eblk[:, j - i1:j - i1 + 1] = err
            blk[:, j - i1 + 1:] -= err @ U[j:j + 1, j + 1:i2]
        W[:, i2:] -= eblk @ U[i1:i2, i2:]
    out = dict(quant=quant, qweight=qweight, error=error, mse_index=None)
    out.update(_collect(params))
    return out


def solve_rtn_groups(weight: torch.Tensor, group_sizes: Sequence[int], bits: Sequence[int], **opts) -> dict:
    """Round to nearest (gptq_py.py:101-124) with group sizes that need not be uniform."""
    W = weight.to(torch.float32).clone()
    quant, qweight = torch.empty_like(W), torch.empty_like(W)
    params, i1 = [], 0
    for n, b in zip(group_sizes, bits):
        i2 = i1 + int(n)
        P = _group_params(W[:, i1:i2], int(b), opts)
        params.append(P)
        qweight[:, i1:i2] = qw = G.quantize(W[:, i1:i2], P['scale'], P['qzero'], P['maxq'])
        quant[:, i1:i2] = G.dequantize(qw, P['scale'], P['qzero'])
        i1 = i2
    out = dict(quant=quant, qweight=qweight, error=None, mse_index=None)
    out.update(_collect(params))
    return out


def assemble_groups(weight_ref: torch.Tensor, solved: dict, U: Optional[torch.Tensor], perm: Optional[torch.Tensor],
                    group_sizes: Sequence[int], bits: Sequence[int], scale_bits: Optional[int] = None) -> dict:
    """gptq_py.py:195-251 for any group sizes, with every optional key: with EXL2 metadata the groups are stored by
    descending (bit width, group size), ties in ascending position (:201-212)."""
    R, C = weight_ref.shape
    dtype = weight_ref.dtype
    sizes = [int(n) for n in group_sizes]
    err = solved['error']
    if err is None:
        metrics = {'gptq_error': 0.0, 'gptq_norm': 0.0}
    else:
        metrics = {'gptq_error': (err ** 2.0).mean().item(),
                   'gptq_norm': ((weight_ref / U.diag()) ** 2.0).mean().item()}
    order = list(range(len(sizes)))
    if scale_bits is not None:
        order.sort(key=lambda g: (int(bits[g]), sizes[g], -g), reverse=True)
    offs = [0]
    for n in sizes:
        offs.append(offs[-1] + n)
    ic = torch.cat([torch.arange(offs[g], offs[g + 1]) for g in order])
    ig = torch.tensor(order)
    p16 = (perm.to(torch.int16) if perm is not None else torch.arange(C, dtype=torch.int16))[ic]
    perm_inv = p16.argsort().to(torch.int16)
    quant = solved['quant'][:, ic][:, perm_inv.to(torch.int32)]
    metrics['l2.0_error'] = ((quant.to(torch.float32) - weight_ref.to(torch.float32)).abs() ** 2.0).mean().item()
    metrics['l2.0_norm'] = (weight_ref.abs() ** 2.0).mean().item()
    has_s = 'sscale' in solved
    meta = {
        'qweight': solved['qweight'].to(torch.uint8)[:, ic],
        'scale': solved['scale'].to(dtype)[:, ig],
        'qzero': solved['qzero'].to(torch.uint8)[:, ig],
        'qscale': solved['qscale'].to(torch.uint8)[:, ig] if has_s else None,
        'sscale': solved['sscale'].to(dtype)[0, ig] if has_s else None,
        'perm_inv': perm_inv,
        'group_sizes': torch.tensor(sizes, dtype=torch.int16)[ig],
        'group_bit_widths': torch.tensor([int(b) for b in bits], dtype=torch.uint8)[ig],
        'scale_bit_width': torch.tensor(scale_bits, dtype=torch.uint8) if scale_bits is not None else None,
    }
    return {'quant_meta': meta, 'metrics': metrics}


def assemble_blocks(weight_ref: torch.Tensor, solved: dict, U: torch.Tensor, perm: Optional[torch.Tensor],
                    group_sizes: Sequence[int], bits: Sequence[int]) -> dict:
    """gptq_py.py:195-251 for solve_blocks (no EXL2 metadata, so the groups keep their order)."""
    R, C = weight_ref.shape
    dtype = weight_ref.dtype
    err = solved['error']
    metrics = {'gptq_error': (err ** 2.0).mean().item(),
               'gptq_norm': ((weight_ref / U.diag()) ** 2.0).mean().item()}
    p16 = perm.to(torch.int16) if perm is not None else torch.arange(C, dtype=torch.int16)
    perm_inv = p16.argsort().to(torch.int16)
    quant = solved['quant'][:, perm_inv.to(torch.int32)]
    metrics['l2.0_error'] = ((quant.to(torch.float32) - weight_ref.to(torch.float32)).abs() ** 2.0).mean().item()
    metrics['l2.0_norm'] = (weight_ref.abs() ** 2.0).mean().item()
    meta = {'qweight': solved['qweight'].to(torch.uint8), 'scale': solved['scale'].to(dtype),
            'qzero': solved['qzero'].to(torch.uint8), 'qscale': None, 'sscale': None, 'perm_inv': perm_inv,
            'group_sizes': torch.tensor([int(n) for n in group_sizes], dtype=torch.int16),
            'group_bit_widths': torch.tensor(list(bits), dtype=torch.uint8), 'scale_bit_width': None}
    return {'quant_meta': meta, 'metrics': metrics}


def group_order(bits: Sequence[int], group_size: int, scale_bits: Optional[int]):
    """gptq_py.py:201-212: with EXL2 metadata the groups are stored by descending bit width (ties: ascending index)."""
    order = list(range(len(bits)))
    if scale_bits is not None:
        order.sort(key=lambda g: (bits[g], group_size, -g), reverse=True)
    cols = torch.cat([torch.arange(g * group_size, (g + 1) * group_size) for g in order])
    return torch.tensor(order), cols


def assemble_general(weight_ref: torch.Tensor, solved: dict, U: Optional[torch.Tensor], perm: Optional[torch.Tensor],
                     group_size: int, bits: Sequence[int], scale_bits: Optional[int] = None) -> dict:
    """gptq_py.py:195-251 with every optional key."""
    R, C = weight_ref.shape
    dtype = weight_ref.dtype
    err = solved['error']
    if err is None:
        metrics = {'gptq_error': 0.0, 'gptq_norm': 0.0}                     # error = tensor(0.)   :124,196-199
    else:
        metrics = {'gptq_error': (err ** 2.0).mean().item(),
                   'gptq_norm': ((weight_ref / U.diag()) ** 2.0).mean().item()}
    ig, ic = group_order(bits, group_size, scale_bits)
    p16 = (perm.to(torch.int16) if perm is not None else torch.arange(C, dtype=torch.int16))[ic]
    perm_inv = p16.argsort().to(torch.int16)
    quant = solved['quant'][:, ic][:, perm_inv.to(torch.int32)]
    metrics['l2.0_error'] = ((quant.to(torch.float32) - weight_ref.to(torch.float32)).abs() ** 2.0).mean().item()
    metrics['l2.0_norm'] = (weight_ref.abs() ** 2.0).mean().item()
    has_s = 'sscale' in solved
    meta = {
        'qweight': solved['qweight'].to(torch.uint8)[:, ic],
        'scale': solved['scale'].to(dtype)[:, ig],
        'qzero': solved['qzero'].to(torch.uint8)[:, ig],
        'qscale': solved['qscale'].to(torch.uint8)[:, ig] if has_s else None,
        'sscale': solved['sscale'].to(dtype)[0, ig] if has_s else None,
        'perm_inv': perm_inv,
        'group_sizes': torch.full((len(bits),), group_size, dtype=torch.int16),
        'group_bit_widths': torch.tensor(list(bits), dtype=torch.uint8)[ig],
        'scale_bit_width': torch.tensor(scale_bits, dtype=torch.uint8) if scale_bits is not None else None,
    }
    return {'quant_meta': meta, 'metrics': metrics}


def reconstruct_dense_general(meta: dict, dtype: torch.dtype) -> torch.Tensor:
    """quant.py:222-279 incl. construct_matrix_2: with EXL2 metadata the scale is qscale^2 * sscale evaluated in int16
    then in the model dtype."""
    if meta.get('sscale') is None:
        return G.reconstruct_dense(meta, dtype)
    q = meta['qweight'].to(torch.int16)
    z = meta['qzero'].to(torch.int16)
    qs = meta['qscale'].to(torch.int16)
    ss = meta['sscale'].to(dtype)
    out = torch.empty(q.shape, dtype=dtype)
    i = 0
    for g, n in enumerate(int(v) for v in meta['group_sizes'].tolist()):
        s = qs[:, g:g + 1] * qs[:, g:g + 1] * ss[g]                         # dequantize2, quant.py:65
        out[:, i:i + n] = (q[:, i:i + n] - z[:, g:g + 1]) * s
        i += n
    return out[:, meta['perm_inv'].to(torch.int32)]
