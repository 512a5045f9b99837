"""the remaining Quantizer options (symmetric grid, MSE grid search, EXL2 quantised scales,
per-group bit widths) through the CUDA path -- gq_find_params, gq_mse_refine, gq_gptq_solve_ex and gptq_quant's three
branches -- against the oracle (oracle/quantizer_oracle.py) and the outputs frozen from the real reference
(tests/golden/q_*.npz).  Integer results are compared code by code; the MSE search compares |e|^2.4 sums whose last bits
depend on the summation order and on powf, so its argmin may move between near-ties: the bound is on the fraction of rows."""
import types

import pytest
import torch

import golden_util as GU
from oracle import gptq_oracle as O
from oracle import quantizer_oracle as Q

pytestmark = pytest.mark.gpu
DEV = torch.device('cuda:0')


def _rand_wt(C, R, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(C, R, generator=g) * 0.02
    w[::7] *= 3.0
    return w


@pytest.mark.parametrize('sym', [False, True])
@pytest.mark.parametrize('R,C,gs,bits', [(64, 256, 128, [4, 3]), (100, 384, 64, [2, 8, 4, 4, 3, 5]), (4096, 512, 128, [4] * 4)])
def test_find_params_is_bit_exact(R, C, gs, bits, sym):
    import gptq as ops
    wt = _rand_wt(C, R, 3 + R)
    wt[5, :] = 0.0
    if R > 8:
        wt[:gs, 7] = 0.0                                              # an all-zero group row: scale = 1e-12
    G = C // gs
    maxq = torch.tensor([2.0 ** b - 1 for b in bits], dtype=torch.float32)
    scale = torch.empty(G, R, device=DEV)
    zero = torch.empty(G, R, device=DEV)
    ops.find_params(wt.to(DEV), gs, maxq.to(DEV), sym, scale, zero)
    for g in range(G):
        P = Q.find_params(wt[g * gs:(g + 1) * gs].t().contiguous(), bits[g], sym=sym)
        assert torch.equal(scale[g].cpu(), P['scale'][:, 0]), g
        assert torch.equal(zero[g].cpu(), P['qzero'][:, 0]), g
    # a sub-range of groups writes only its own rows
    s2 = torch.full_like(scale, -1.0)
    z2 = torch.full_like(zero, -1.0)
    ops.find_params(wt.to(DEV), gs, maxq.to(DEV), sym, s2, z2, g0=1, n_groups=1)
    assert torch.equal(s2[1], scale[1]) and torch.equal(z2[1], zero[1])
    assert bool((s2[0] == -1).all()) and (G < 3 or bool((s2[2] == -1).all()))


@pytest.mark.parametrize('norm', [2.4, 2.0])
@pytest.mark.parametrize('sym,bits', [(False, 4), (True, 3), (False, 2)])
def test_mse_refine_matches_oracle(sym, bits, norm):
    import gptq as ops
    R, gs, G = 96, 128, 3
    wt = _rand_wt(G * gs, R, 11 + bits)
    maxq = torch.full([G], 2.0 ** bits - 1)
    scale = torch.empty(G, R, device=DEV)
    zero = torch.empty(G, R, device=DEV)
    ops.find_params(wt.to(DEV), gs, maxq.to(DEV), sym, scale, zero)
    s0 = scale.clone()
    grid = Q.shrink_grid(0.8, 100)
    idx = ops.mse_refine(wt.to(DEV), gs, maxq.to(DEV), grid.to(DEV), norm, scale, zero, want_index=True)
    same = 0
    for g in range(G):
        x = wt[g * gs:(g + 1) * gs].t().contiguous()
        P = Q.find_params(x, bits, sym=sym)
        Q.mse_refine(P, x, 0.8, 100, norm)
        want = P['mse_index'][:, 0]
        got = idx[g].cpu().long()
        same += int((want == got).sum())
        # the stored scale is always scale0 * grid[chosen index], bit for bit
        assert torch.equal(scale[g].cpu(), s0[g].cpu() * grid[got])
        # where the index differs the two candidates are near-ties of the reference's own error sums
        for r in (want != got).nonzero().flatten().tolist():
            e = []
            for i in (int(want[r]), int(got[r])):
                s = s0[g, r].cpu() * grid[i]
                q = O.dequantize(O.quantize(x[r], s, P['qzero'][r], P['maxq']), s, P['qzero'][r])
                e.append(float((q - x[r]).abs().pow(norm).sum()))
            assert abs(e[0] - e[1]) <= 1e-4 * max(e), (g, r, e)
    assert same >= 0.97 * G * R, same / (G * R)


def _fake_hook(g):
    if g['U'] is None:
        return None
    return types.SimpleNamespace(hessian_inv=g['U'].to(DEV), perm=(g['perm'].to(DEV) if g['perm'] is not None else None),
                                 hessian=None)


def _run(g, W, use_kernel):
    from gptq_py import gptq_quant
    G = g['C'] // 128
    gs = torch.full([G], 128, dtype=torch.int32, device=DEV)
    bw = torch.tensor(g['bits'], dtype=torch.int32, device=DEV)
    sb = torch.tensor(g['scale_bits']) if g['scale_bits'] is not None else None
    return gptq_quant(weight=W.to(DEV), hessian_hook=_fake_hook(g), group_sizes=gs, group_bit_widths=bw, scale_bit_width=sb,
                      gptq_use_kernel=use_kernel, gptq_block_sizes=gs, quant_symmetric=g['sym'], quant_mse=g['mse'],
                      quant_max_shrink=.8, quant_n_grid=100, quant_norm=2.4, quant_use_kernel=False)


def _row_agreement(a, b):
    """fraction of rows of two (R, ...) tensors that are identical"""
    return float((a.reshape(a.shape[0], -1) == b.reshape(b.shape[0], -1)).all(dim=1).float().mean())


@pytest.mark.parametrize('name', GU.qnames())
def test_gptq_quant_options_match_reference_outputs(name):
    """gptq_quant on the GPU (fp32 / round-to-nearest branches) vs the real reference's frozen outputs."""
    from quant import reconstruct_nn_linear
    g = GU.loadq(name)
    res = _run(g, g['W'], use_kernel=False)
    m = res['quant_meta']
    assert torch.equal(m['perm_inv'], g['perm_inv'])
    assert torch.equal(m['group_bit_widths'], g['group_bit_widths'])
    assert m['qweight'].dtype == torch.uint8 and m['scale'].dtype == g['dtype']
    exl2 = g['scale_bits'] is not None
    if exl2:
        assert m['qscale'].dtype == torch.uint8 and m['sscale'].dtype == g['dtype'] and int(m['scale_bit_width']) == g['scale_bits']
        assert m['sscale'].shape == g['sscale'].shape
    else:
        assert m['qscale'] is None and m['sscale'] is None and m['scale_bit_width'] is None
    # without the MSE search everything is determined up to the solver's cross-block fp32 sums
    codes = O.code_match(m['qweight'], g['qweight'])
    rows = _row_agreement(m['scale'].float(), g['scale'])
    if not g['mse']:
        assert codes >= 0.999, codes
        assert torch.equal(m['qzero'], g['qzero']) or (m['qzero'].int() - g['qzero'].int()).abs().max() <= 1
        assert torch.allclose(m['scale'].float(), g['scale'], rtol=2e-3)
        if exl2:
            assert _row_agreement(m['qscale'], g['qscale']) >= 0.98
            assert torch.allclose(m['sscale'].float(), g['sscale'], rtol=2e-3)
    else:
        # a different shrink factor in a row changes that row's codes of the group; rows with the same factor agree
        # (measured on B200: all eight cases reproduce the reference's codes and scales exactly)
        assert rows >= 0.98, rows
        assert codes >= 0.985, codes
    tol = 2e-2
    for k in ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'):
        assert res['metrics'][k] == pytest.approx(g['metrics'][k], rel=tol, abs=1e-30), k
    W_hat = reconstruct_nn_linear(m, dtype=g['dtype'], device=DEV).weight.data.float().cpu()
    # the quantised matrix is as good as the reference's: same distance to W within a few percent
    e_ref = (g['W_hat'] - g['W'].float()).norm()
    e_got = (W_hat - g['W'].float()).norm()
    assert abs(float(e_got - e_ref)) <= 3e-2 * float(e_ref), (float(e_got), float(e_ref))
    # and the dense reconstruction reproduces this build's own metadata exactly as the oracle's formula does
    want = Q.reconstruct_dense_general({k: (v.cpu() if torch.is_tensor(v) else v) for k, v in m.items()}, g['dtype'])
    assert torch.equal(W_hat, want.float())


@pytest.mark.parametrize('sym,sbits,mse,bits', [(True, None, False, [4, 4, 4]), (False, None, True, [4, 3, 4]),
                                                (False, 4, False, [3, 4, 2]), (False, 4, True, [4, 4, 4])])
def test_kernel_branch_options_vs_oracle(sym, sbits, mse, bits):
    """gptq_use_kernel=True with every option vs the oracle's restatement of the reference's kernel branch."""
    R, C = 64, 384
    W, X = O.synth_linear(R, C, 4, 128, dtype=torch.float16, seed=901)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    g = dict(C=C, bits=bits, scale_bits=sbits, sym=sym, mse=mse, U=inv['U'], perm=inv['perm'])
    res = _run(g, W, use_kernel=True)
    sol = Q.solve_general(W, inv['U'], inv['perm'], 128, bits, 'kernel', sym=sym, scale_bits=sbits, mse=mse)
    want = Q.assemble_general(W, sol, inv['U'], inv['perm'], 128, bits, sbits)
    m, w = res['quant_meta'], want['quant_meta']
    assert torch.equal(m['perm_inv'], w['perm_inv']) and torch.equal(m['group_bit_widths'], w['group_bit_widths'])
    rows = _row_agreement(m['scale'].float(), w['scale'].float())
    codes = O.code_match(m['qweight'], w['qweight'])
    if mse:
        assert rows >= 0.90 and codes >= 0.93, (rows, codes)
    else:
        assert rows == 1.0 and torch.equal(m['qzero'], w['qzero'])           # static parameters: bit-exact
        assert codes >= 0.999, codes
        if sbits is not None:
            assert torch.equal(m['qscale'], w['qscale']) and torch.equal(m['sscale'], w['sscale'])
    for k in ('gptq_error', 'l2.0_error'):
        assert res['metrics'][k] == pytest.approx(want['metrics'][k], rel=6e-2 if mse else 2e-2), k


def test_solve_ex_mode4_equals_mode0_on_asymmetric_parameters():
    """Mode 4 with the parameters mode 0 computed itself reproduces mode 0 bit for bit (same kernel arithmetic), group by
    group through the [a, b) interface."""
    import gptq as ops
    R, C, gs = 128, 512, 128
    W, X = O.synth_linear(R, C, 4, 128, dtype=torch.float16, seed=77)
    H, n = O.hessian_accumulate(None, 0, X)
    U = O.hessian_invert(H, n, 1e-2, False)['U'].to(DEV)
    maxq = [15.0, 7.0, 15.0, 3.0]
    wt = W.float().t().contiguous().to(DEV)
    ops.set_solve_super(1)
    try:
        s0 = torch.empty(C // gs, R, device=DEV)
        z0 = torch.empty_like(s0)
        q0, c0, e0 = ops.gptq_solve_ex(wt.clone(), U, s0, z0, gs, maxq, mode=0)
        wt4 = wt.clone()
        out = None
        for g in range(C // gs):
            out = ops.gptq_solve_ex(wt4, U, s0, z0, gs, maxq, mode=4, a=g * gs, b=(g + 1) * gs, out=out)
    finally:
        ops.set_solve_super(8)
    assert torch.equal(out[1], c0) and torch.equal(out[0], q0) and torch.equal(out[2], e0)


@pytest.mark.parametrize('name', GU.bnames())
def test_gptq_quant_blocks_and_nonuniform_groups(name):
    """gptq_block_sizes != group_sizes (nested either way or not at all), non-uniform group sizes, with and without the MSE
    search over the block's columns / EXL2 scales, on the fp32 and round-to-nearest branches, against the outputs of the real
    reference (tests/golden/b_*.npz, gptq_py.py:101-124,161-193).  Blocks smaller than groups repeat the reference's own
    blocking; otherwise the same updates are applied in another order (fp32 reordering: rounding ties may move)."""
    from gptq_py import gptq_quant
    from quant import reconstruct_nn_linear
    g = GU.loadb(name)
    hook = None
    if not g['rtn']:
        hook = types.SimpleNamespace(hessian_inv=g['U'].to(DEV), perm=(g['perm'].to(DEV) if g['perm'] is not None else None),
                                     perm_inv=(torch.argsort(g['perm']).to(DEV) if g['perm'] is not None else None), hessian=None)
    res = gptq_quant(weight=g['W'].to(DEV), hessian_hook=hook,
                     group_sizes=torch.tensor(g['group_sizes'], dtype=torch.int32),
                     group_bit_widths=torch.tensor(g['bits'], dtype=torch.int32),
                     scale_bit_width=(torch.tensor(g['scale_bits']) if g['scale_bits'] is not None else None),
                     gptq_use_kernel=False,
                     gptq_block_sizes=(torch.tensor(g['block_sizes'], dtype=torch.int32) if not g['rtn'] else None),
                     quant_symmetric=g['sym'], quant_mse=g['mse'])
    m = res['quant_meta']
    assert torch.equal(m['perm_inv'], g['perm_inv'])
    if g['group_sizes_out'] is not None:
        assert torch.equal(m['group_sizes'], g['group_sizes_out'])
        assert torch.equal(m['group_bit_widths'], g['group_bit_widths_out'])
    match = O.code_match(m['qweight'], g['qweight'])
    rows = _row_agreement(m['scale'].float(), g['scale'])
    if g['rtn'] and not g['mse']:
        assert match == 1.0 and torch.equal(m['qzero'], g['qzero']) and rows == 1.0       # nothing propagates: bit-exact
    elif not g['mse']:
        assert match >= 0.999, match
        assert (m['qzero'].float() - g['qzero'].float()).abs().max().item() <= 1
        assert (m['qzero'] == g['qzero']).float().mean().item() >= 0.995
        assert torch.allclose(m['scale'].float(), g['scale'], rtol=2e-3, atol=0)
    else:
        # a different shrink factor in a row (near-ties of |e|^2.4 sums) changes that row's codes of the group
        assert rows >= 0.95, rows
        assert match >= 0.97, match
    if g['scale_bits'] is not None:
        assert m['qscale'].dtype == torch.uint8 and m['sscale'].shape == g['sscale'].shape
        assert torch.allclose(m['sscale'].float(), g['sscale'], rtol=2e-2)
    for k in ('gptq_error', 'l2.0_error'):
        assert res['metrics'][k] == pytest.approx(g['metrics'][k], rel=(5e-2 if g['mse'] else 1e-2), abs=1e-30), k
    # the dense reconstruction of this build's own metadata follows the reference's formula for any group sizes
    W_hat = reconstruct_nn_linear(m, dtype=g['dtype'], device=DEV).weight.data.float().cpu()
    want = Q.reconstruct_dense_general({k: (v.cpu() if torch.is_tensor(v) else v) for k, v in m.items()}, g['dtype'])
    assert torch.equal(W_hat, want.float())
    e_ref = (g['W_hat'] - g['W'].float()).norm()
    e_got = (W_hat - g['W'].float()).norm()
    assert abs(float(e_got - e_ref)) <= 3e-2 * float(e_ref), (float(e_got), float(e_ref))


def test_kernel_branch_with_nonuniform_and_wide_groups():
    """gptq_use_kernel=True for group sizes the one-call solver does not take (non-uniform, 256): the reference's own
    per-group gptq_quantize_range loop (gptq_py.py:135-139), against the oracle's restatement of quantize_range."""
    from gptq_py import gptq_quant
    R, C = 64, 512
    W, X = O.synth_linear(R, C, 4, 128, dtype=torch.float16, seed=4242)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    hook = types.SimpleNamespace(hessian_inv=inv['U'].to(DEV), perm=inv['perm'].to(DEV), perm_inv=inv['perm_inv'].to(DEV), hessian=None)
    for sizes in ([256, 256], [64, 192, 96, 160]):
        bits = [4, 3, 4, 4][:len(sizes)]
        res = gptq_quant(weight=W.to(DEV), hessian_hook=hook, group_sizes=torch.tensor(sizes, dtype=torch.int32),
                         group_bit_widths=torch.tensor(bits, dtype=torch.int32), gptq_use_kernel=True, quant_mse=False)
        Wp = W.float()[:, inv['perm']]
        wt = Wp.t().contiguous()
        quant_t, err_t = torch.empty_like(wt), torch.empty_like(wt)
        q_t = torch.empty_like(wt, dtype=torch.int16)
        c0, scales = 0, []
        for sz, b in zip(sizes, bits):
            P = Q.find_params(Wp[:, c0:c0 + sz], b)
            scales.append(P['scale'])
            O.quantize_range_ref(quant_t, P['scale'][:, 0].contiguous(), q_t, P['qzero'][:, 0].contiguous(), float(P['maxq']),
                                 inv['U'], wt, err_t, c0, c0 + sz)
            c0 += sz
        m = res['quant_meta']
        assert O.code_match(m['qweight'], q_t.t()) >= 0.999
        assert torch.equal(m['scale'].float(), torch.cat(scales, dim=1).to(torch.float16).float())   # static parameters: bit-exact
        assert m['group_sizes'].tolist() == sizes


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16, torch.float32])
def test_finalize_quant_sums_per_row_segment(dtype):
    """gq_finalize_quant: the four sums per row segment and the transposed scale / zero tables against plain torch on the same
    arrays (gptq_py.py:196-199, 231-237 semantics: W unpermuted against the permuted-order diag(U); squares rounded to the
    weight dtype for the l2 norm)."""
    import gptq
    R, C, G = 416, 384, 3
    rows = [160, 64, 192]
    g = torch.Generator(device='cuda').manual_seed(11)
    err_t = torch.randn(C, R, device='cuda', generator=g)
    quant_t = torch.randn(C, R, device='cuda', generator=g) * 0.02
    orig_t = (torch.randn(C, R, device='cuda', generator=g) * 0.02).to(dtype)
    scale_gr = torch.rand(G, R, device='cuda', generator=g) + 0.5
    qzero_gr = torch.randint(0, 16, (G, R), device='cuda', generator=g).float()
    perm = torch.randperm(C, device='cuda', generator=g)
    U = torch.triu(torch.rand(C, C, device='cuda', generator=g) + 0.5)
    for segs, pm, u in ((rows, perm, U), (None, None, U), (rows, perm, None)):
        stats, scale, qzero = gptq.finalize_quant(err_t, quant_t, orig_t, scale_gr, qzero_gr, pm, u, segs)
        torch.cuda.synchronize()
        assert torch.equal(scale, scale_gr.t().to(dtype).contiguous()) and torch.equal(qzero, qzero_gr.t().to(torch.uint8).contiguous())
        lo = 0
        for si, n in enumerate(segs or [R]):
            sl = slice(lo, lo + n)
            o = orig_t[:, sl].double()
            want = [float((err_t[:, sl].double() ** 2).sum()), float(((quant_t[:, sl].double() - o) ** 2).sum())]
            if u is not None:
                d = u.diagonal().double()[pm] if pm is not None else u.diagonal().double()
                want.append(float(((o ** 2).sum(dim=1) / d ** 2).sum()))
            else:
                want.append(0.0)
            want.append(float((orig_t[:, sl] * orig_t[:, sl]).double().sum()))
            got = stats[si].tolist()
            for a, b in zip(got, want):
                assert a == pytest.approx(b, rel=2e-6, abs=1e-12), (si, got, want)
            lo += n
    with pytest.raises(gptq.GptqB200Error):
        gptq.finalize_quant(err_t, quant_t, orig_t, scale_gr, qzero_gr, perm, U, [100, 316 + 1])
