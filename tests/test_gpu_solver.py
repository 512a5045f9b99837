"""K3/K4 parity: block solver + lazy-batch update vs the oracle restatements of
gptq_py.py:158-193 (mode 0) and gptq_cpp.cpp:33-84 + gptq_cpp_kernel.cu (mode 1)."""
import pytest
import torch

import golden_util as GU
from oracle import gptq_oracle as O

pytestmark = pytest.mark.gpu


def _setup(R, C, seed, n_seq=8, seqlen=256):
    W, X = O.synth_linear(R, C, n_seq, seqlen, seed=seed)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    return W, inv


@pytest.mark.parametrize('R,C,group,bits', [(64, 256, 128, 4), (100, 384, 128, 3), (256, 512, 64, 4),
                                            (96, 256, 32, 4), (512, 1024, 128, 4)])
def test_solver_pure_mode_vs_oracle(R, C, group, bits):
    """Same U injected on both sides.  First block is bit-exact (no GEMM involved); overall
    >= 99.9% of codes equal, mismatches only where the oracle sits at a rounding tie."""
    import gptq
    W, inv = _setup(R, C, seed=R + C)
    sol = O.gptq_solve_pure(W, inv['U'], inv['perm'], group, bits)
    Wt = gptq.prep_weight_t(W.cuda().contiguous(), inv['perm'].cuda())
    assert torch.equal(Wt.cpu(), W.float()[:, inv['perm']].t())
    quant_t, codes_t, err_t, s_gr, z_gr = gptq.gptq_solve(Wt, inv['U'].cuda().contiguous(), group, float(2 ** bits - 1))
    torch.cuda.synchronize()
    codes = codes_t.t().cpu().float()
    # block 0: exact
    assert torch.equal(codes[:, :group], sol['qweight'][:, :group])
    assert torch.equal(s_gr[0].cpu(), sol['scale'][:, 0])
    assert torch.equal(z_gr[0].cpu(), sol['qzero'][:, 0])
    assert torch.equal(quant_t.t().cpu()[:, :group], sol['quant'][:, :group])
    assert torch.equal(err_t.t().cpu()[:, :group], sol['error'][:, :group])
    match = O.code_match(codes, sol['qweight'])
    assert match >= 0.999, match
    # a flipped code (allowed: < 0.1%) changes the compensation of the rest of its row, so scales of later
    # groups may differ in those rows; everywhere else they agree to fp32 reordering level
    close = torch.isclose(s_gr.t().cpu(), sol['scale'], rtol=1e-4)
    assert close.float().mean().item() >= 0.98
    # reconstruction error within 1% of the oracle's
    e_gpu = O.recon_error(W, quant_t.t().cpu(), inv['H'], inv['perm'])
    e_ora = O.recon_error(W, sol['quant'], inv['H'], inv['perm'])
    assert abs(e_gpu - e_ora) <= 0.01 * e_ora, (e_gpu, e_ora)


@pytest.mark.parametrize('R,C,bits', [(64, 256, 4), (128, 512, 3)])
def test_quantize_range_compat_vs_oracle(R, C, bits):
    """gptq.gptq_quantize_range with the reference's argument list on the transposed layout."""
    import gptq
    W, inv = _setup(R, C, seed=7 * R + C)
    sol = O.gptq_solve_kernelpath(W, inv['U'], inv['perm'], 128, bits)
    scale, qzero, maxq = O.static_group_params(W, inv['perm'], 128, bits)
    dev = 'cuda'
    wt = W.float()[:, inv['perm']].t().contiguous().to(dev)
    quant_t = torch.empty_like(wt)
    err_t = torch.empty_like(wt)
    q_t = torch.empty_like(wt, dtype=torch.int16)
    U = inv['U'].to(dev).contiguous()
    for g, i1 in enumerate(range(0, C, 128)):
        gptq.gptq_quantize_range(quant_t, scale[:, g].contiguous().to(dev), q_t, qzero[:, g].contiguous().to(dev),
                                 maxq, U, wt, err_t, i1, i1 + 128)
    torch.cuda.synchronize()
    q = q_t.t().cpu().float()
    assert torch.equal(q[:, :128], sol['qweight'][:, :128])
    assert torch.equal(quant_t.t().cpu()[:, :128], sol['quant'][:, :128])
    assert torch.equal(err_t.t().cpu()[:, :128], sol['error'][:, :128])
    assert O.code_match(q, sol['qweight']) >= 0.999
    # meta out_q is skipped like the reference (gptq_cpp.cpp:58)
    wt2 = W.float()[:, inv['perm']].t().contiguous().to(dev)
    gptq.gptq_quantize_range(quant_t, scale[:, 0].contiguous().to(dev), torch.empty(0, device='meta'),
                             qzero[:, 0].contiguous().to(dev), maxq, U, wt2, err_t, 0, 128)
    with pytest.raises(RuntimeError):
        gptq.gptq_quantize_range(quant_t, scale[:, 0].contiguous().to(dev), q_t, qzero[:, 0].contiguous().to(dev),
                                 maxq, U[:, :-4].contiguous(), wt2, err_t, 0, 128)


@pytest.mark.parametrize('R,C,group', [(64, 256, 128), (128, 512, 64), (256, 1024, 128), (32, 128, 32)])
def test_solve_static_equals_per_group_range_calls(R, C, group):
    """gq_gptq_solve_static = the reference's `for g: gptq_quantize_range(...)` loop (gptq_py.py:135-139), bit for bit."""
    import gptq
    W, inv = _setup(R, C, seed=3 * R + C)
    scale, qzero, maxq = O.static_group_params(W, inv['perm'], group, 4)
    dev = 'cuda'
    U = inv['U'].to(dev).contiguous()
    wt = W.float()[:, inv['perm']].t().contiguous().to(dev)
    quant_t = torch.empty_like(wt)
    err_t = torch.empty_like(wt)
    q_t = torch.empty_like(wt, dtype=torch.int16)
    for g, i1 in enumerate(range(0, C, group)):
        gptq.gptq_quantize_range(quant_t, scale[:, g].contiguous().to(dev), q_t, qzero[:, g].contiguous().to(dev),
                                 maxq, U, wt, err_t, i1, i1 + group)
    # one full trailing update per block (the reference's schedule): bit for bit
    old = gptq.set_solve_super(1)
    try:
        wt2 = W.float()[:, inv['perm']].t().contiguous().to(dev)
        quant2, q2, err2 = gptq.gptq_solve_static(wt2, U, scale.t().contiguous().to(dev), qzero.t().contiguous().to(dev), group, maxq)
        torch.cuda.synchronize()
    finally:
        gptq.set_solve_super(old)
    assert torch.equal(q_t, q2) and torch.equal(quant_t, quant2) and torch.equal(err_t, err2) and torch.equal(wt, wt2)
    # default two-level schedule: same arithmetic, the far updates summed in a different order
    wt3 = W.float()[:, inv['perm']].t().contiguous().to(dev)
    quant3, q3, err3 = gptq.gptq_solve_static(wt3, U, scale.t().contiguous().to(dev), qzero.t().contiguous().to(dev), group, maxq)
    torch.cuda.synchronize()
    assert torch.equal(q_t[:group], q3[:group])
    assert (q_t == q3).float().mean().item() >= 0.999
