"""A/B parity on the B200 against the REFERENCE's own CUDA extension (`gptq_c`), compiled by
oracle/build_ref.py from /root/reference/package/gptq into oracle/_ref/gptq_c.so (git-ignored, travels to the GPU box).

  a2  accumulate_hessian_{fp16,bf16}_fp32  (CUTLASS sm80 GEMM, accumulate_hessian_kernel.cu:10-84)
  a6  gptq_quantize_range                  (gptq_cpp.cpp:33-84 + gptq_cpp_kernel.cu:21-158)

These pin the two compat entry points against the thing they replace, on identical device buffers.  They skip (not
fail) when oracle/_ref was not built, e.g. a checkout without /root/reference.
"""
import pytest
import torch

import ref_ext
from oracle import gptq_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def ref():
    m = ref_ext.load()
    if m is None:
        pytest.skip('oracle/_ref/gptq_c.so not built (needs /root/reference at build time)')
    return m


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('D,T', [(512, 256), (1024, 4096), (4096, 2048), (1408, 520)])
def test_accumulate_hessian_vs_reference_cutlass(ref, D, T, dtype):
    """Same call as package/tests/test.py:10-27 (randn H0, asymmetric), reference kernel vs ours, both RMW the full H."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(D + T)
    X = torch.randn(T, D, device='cuda', generator=g).to(dtype)
    H0 = torch.randn(D, D, device='cuda', generator=g)
    H_ref, H_our = H0.clone(), H0.clone()
    fn = ref.accumulate_hessian_fp16_fp32 if dtype == torch.float16 else ref.accumulate_hessian_bf16_fp32
    fn(H_ref, X)
    gptq.accumulate_hessian(H_our, X)
    torch.cuda.synchronize()
    err = O.rel_fro(H_our.cpu(), H_ref.cpu())
    assert err <= 1e-3, err                 # north_star bound; both are 16-bit x 16-bit -> fp32, so ~1e-7 is expected
    assert err <= 1e-5, err
    # the reference's own check: against fp32 torch (package/tests/test.py:22-27)
    want = H0 + X.float().t() @ X.float()
    assert O.rel_fro(H_our.cpu(), want.cpu()) <= O.rel_fro(H_ref.cpu(), want.cpu()) * 2 + 1e-6


@pytest.mark.parametrize('R,C,bits', [(64, 256, 4), (256, 512, 4), (128, 384, 3), (1024, 1024, 4)])
def test_quantize_range_vs_reference_kernels(ref, R, C, bits):
    """The reference's per-column launches + addmm_ against our fused block solver + 3xTF32 update, group by group,
    exactly as gptq_py.py:129-147 drives them."""
    import gptq
    W, X = O.synth_linear(R, C, 8, 256, seed=3 * R + C)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    scale, qzero, maxq = O.static_group_params(W, inv['perm'], 128, bits)
    dev = 'cuda'
    U = inv['U'].to(dev).contiguous()
    torch.backends.cuda.matmul.allow_tf32 = False            # main.py:12-13
    out = {}
    for who in ('ref', 'our'):
        wt = W.float()[:, inv['perm']].t().contiguous().to(dev)
        quant_t, err_t = torch.empty_like(wt), torch.empty_like(wt)
        q_t = torch.zeros_like(wt, dtype=torch.int16)
        for g_, i1 in enumerate(range(0, C, 128)):
            args = (quant_t, scale[:, g_].contiguous().to(dev), q_t, qzero[:, g_].contiguous().to(dev), maxq, U, wt,
                    err_t, i1, i1 + 128)
            if who == 'ref':
                ref.gptq_quantize_range(*args)
            else:
                gptq.gptq_quantize_range(*args)
        torch.cuda.synchronize()
        out[who] = (q_t.cpu(), quant_t.cpu(), err_t.cpu(), wt.cpu())
    q_r, quant_r, err_r, w_r = out['ref']
    q_o, quant_o, err_o, w_o = out['our']
    # first 128 columns involve no GEMM: bit-exact codes, dequantised values and errors
    assert torch.equal(q_o[:128], q_r[:128])
    assert torch.equal(quant_o[:128], quant_r[:128])
    assert torch.equal(err_o[:128], err_r[:128])
    match = (q_o == q_r).float().mean().item()
    assert match >= 0.999, match
    # a flipped code (allowed: < 0.1%) changes the compensation of the rest of its row, so errors and weights are
    # compared on the rows without any flip; there they agree to fp32 reordering noise of the trailing GEMM
    rows_ok = (q_o == q_r).all(dim=0)
    assert rows_ok.float().mean().item() >= 0.5
    assert (err_o[:, rows_ok] - err_r[:, rows_ok]).abs().max() <= 1e-3 * err_r.abs().max()
    assert (w_o[:, rows_ok] - w_r[:, rows_ok]).abs().max() <= 1e-4 * w_r.abs().max()
