"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol
declared in include/gptq_b200.h, and the product refuses CPU tensors (no fallback)."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, 'include', 'gptq_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b((?:gq_|accumulate_hessian_)\w+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    from gptq import _cabi
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(_cabi.lib, n), n
        assert n in _cabi.SIGNATURES, f'{n} declared in the header but not bound'
    assert set(_cabi.SIGNATURES) == set(names)
    assert _cabi.lib.gq_abi_version() == 1
    assert _cabi.lib.gq_error_string(2).decode().startswith('unsupported shape')
    assert _cabi.lib.gq_invert_workspace_floats(128) >= 128 * 128 * 3 // 2


def test_reference_surface_is_importable():
    import gptq
    import gptq_py
    from gptq_triton.accumulate_hessian import accumulate_hessian
    import inspect
    assert list(inspect.signature(gptq.accumulate_hessian).parameters) == ['mat_hessian', 'mat_input']
    assert list(inspect.signature(gptq.gptq_quantize_range).parameters) == [
        'quant', 'scale', 'out_q', 'qzero', 'maxq', 'hessian_inv', 'weights', 'error', 'a', 'b']
    assert list(inspect.signature(accumulate_hessian).parameters) == [
        'mat_hessian', 'mat_input', 'save_lower_only', 'compute_lower_only']
    sig = inspect.signature(gptq_py.gptq_quant)
    ref_params = ['weight', 'hessian_hook', 'group_sizes', 'group_bit_widths', 'scale_bit_width',
                  'gptq_use_kernel', 'gptq_block_sizes', 'quant_symmetric', 'quant_mse',
                  'quant_max_shrink', 'quant_n_grid', 'quant_norm', 'quant_use_kernel', 'save_device']
    assert list(sig.parameters)[:len(ref_params)] == ref_params
    # anything beyond the reference's parameters is keyword-only with a default that keeps the reference behaviour
    for extra in list(sig.parameters.values())[len(ref_params):]:
        assert extra.kind is inspect.Parameter.KEYWORD_ONLY and extra.default in (False, None)
    assert sig.parameters['gptq_use_kernel'].default is True and sig.parameters['quant_mse'].default is True
    h = gptq_py.HessianHook()
    assert (h.hessian, h.n_samples, h.hessian_inv, h.perm, h.perm_inv) == (None, 0, None, None, None)


def test_no_cpu_fallback():
    import gptq
    with pytest.raises(gptq.GptqB200Error):
        gptq.hessian_syrk(torch.zeros(64, 64), torch.zeros(8, 64, dtype=torch.float16))
    with pytest.raises(gptq.GptqB200Error):
        gptq.pack_int4(torch.zeros(4, 8, dtype=torch.uint8))
    with pytest.raises(NotImplementedError):
        gptq.accumulate_hessian(torch.zeros(64, 64), torch.zeros(8, 64))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'gptq-v2_b200')
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(dp, fn)).read()
                assert 'oracle' not in txt.replace('the oracle', '').replace('parity oracle', ''), os.path.join(dp, fn)
