"""`gptq_c`: the reference's native module (/root/reference/package/gptq/pybind.cpp:31-35) built by this repository
over libgptq_b200.so (gptq-v2_b200/native/gptq_c.cpp).  The wrappers below are the reference's own
package/gptq/__init__.py:6-28 restated (dtype dispatch, argument order), so the GPU test drives the module exactly the
way the reference's gptq_py.py would after `import gptq_c`."""
import importlib.util
import os

import pytest
import torch

from oracle import gptq_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, 'gptq-v2_b200', 'lib', 'gptq_c.so')


def _load():
    spec = importlib.util.spec_from_file_location('gptq_c', SO)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def accumulate_hessian(gptq_c, mat_hessian, mat_input):            # package/gptq/__init__.py:6-13
    if mat_input.dtype == torch.float16:
        gptq_c.accumulate_hessian_fp16_fp32(mat_hessian, mat_input)
    elif mat_input.dtype == torch.bfloat16:
        gptq_c.accumulate_hessian_bf16_fp32(mat_hessian, mat_input)
    else:
        raise NotImplementedError


def test_module_exports_the_reference_names():
    assert os.path.exists(SO), 'run python __graft_entry__.py (or gptq-v2_b200/native/build_gptq_c.py)'
    m = _load()
    assert {n for n in dir(m) if not n.startswith('_')} == {'accumulate_hessian_fp16_fp32', 'accumulate_hessian_bf16_fp32',
                                                              'gptq_quantize_range'}
    with pytest.raises(RuntimeError):                               # no CPU path: loud failure, like the ctypes binding
        m.accumulate_hessian_fp16_fp32(torch.zeros(8, 8), torch.zeros(4, 8, dtype=torch.float16))


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
def test_gptq_c_accumulate_hessian(dtype):
    m = _load()
    D, T = 640, 777
    g = torch.Generator(device='cuda').manual_seed(1)
    X = torch.randn(T, D, device='cuda', generator=g).to(dtype)
    H = torch.randn(D, D, device='cuda', generator=g)
    want = H.double() + X.double().t() @ X.double()
    accumulate_hessian(m, H, X)
    assert float((H.double() - want).norm() / want.norm()) <= 1e-5
    with pytest.raises(RuntimeError):
        m.accumulate_hessian_fp16_fp32(H, X.float())                # dtype mismatch
    with pytest.raises(RuntimeError):
        m.accumulate_hessian_fp16_fp32(H[:, :-8].contiguous(), X.half())   # 'H shape mismatch'


@pytest.mark.gpu
def test_gptq_c_quantize_range_equals_ctypes_binding():
    """Same library underneath: the pybind module and the ctypes package give bit-identical results, and both follow the
    oracle's restatement of quantize_range (gptq_cpp.cpp:33-84) on the first block."""
    import gptq
    m = _load()
    R, C = 64, 384
    W, X = O.synth_linear(R, C, 4, 128, seed=12)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    scale, qzero, maxq = O.static_group_params(W, inv['perm'], 128, 4)
    U = inv['U'].cuda().contiguous()
    outs = []
    for fn in (m.gptq_quantize_range, gptq.gptq_quantize_range):
        wt = W.float()[:, inv['perm']].t().contiguous().cuda()
        quant_t, err_t = torch.zeros_like(wt), torch.zeros_like(wt)
        q_t = torch.zeros_like(wt, dtype=torch.int16)
        for gi, i1 in enumerate(range(0, C, 128)):
            fn(quant_t, scale[:, gi].contiguous().cuda(), q_t, qzero[:, gi].contiguous().cuda(), maxq, U, wt, err_t, i1, i1 + 128)
        torch.cuda.synchronize()
        outs.append((quant_t, q_t, err_t, wt))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    sol = O.gptq_solve_kernelpath(W, inv['U'], inv['perm'], 128, 4)
    assert torch.equal(outs[0][1].t().cpu().float()[:, :128], sol['qweight'][:, :128])
    assert O.code_match(outs[0][1].t().cpu().float(), sol['qweight']) >= 0.999
    # meta out_q is skipped like the reference (gptq_cpp.cpp:58); a wrong Hessian shape raises
    wt = W.float()[:, inv['perm']].t().contiguous().cuda()
    m.gptq_quantize_range(outs[0][0], scale[:, 0].contiguous().cuda(), torch.empty(0, device='meta'), qzero[:, 0].contiguous().cuda(),
                          maxq, U, wt, outs[0][2], 0, 128)
    with pytest.raises(RuntimeError):
        m.gptq_quantize_range(outs[0][0], scale[:, 0].contiguous().cuda(), outs[0][1], qzero[:, 0].contiguous().cuda(), maxq,
                              U[:, :-4].contiguous(), wt, outs[0][2], 0, 128)
