"""End-to-end drop-in parity: this build's HessianHook + gptq_quant (CUDA) vs outputs frozen
from the real reference (tests/golden) and vs the oracle on larger synthetic linears, judged
by BASELINE.json's criteria: H rel-Frobenius <= 1e-3, >= 99.9% codes equal, reconstruction
error within 1%, dequant-GEMM rel err <= 1e-2."""
import pytest
import torch

import golden_util as GU
from oracle import gptq_oracle as O

pytestmark = pytest.mark.gpu


def _run_gpu(W, X, bits, act_order, use_kernel, batch=2):
    from gptq_py import HessianHook, gptq_quant
    hook = HessianHook()
    Xc = X.cuda()
    for i in range(0, X.shape[0], batch):
        hook.add_batch(Xc[i:i + batch])
    hook.invert(1e-2, act_order)
    C = W.shape[1]
    gs = torch.full([C // 128], 128, dtype=torch.int32, device='cuda')
    bw = torch.full([C // 128], bits, dtype=torch.int32, device='cuda')
    res = gptq_quant(weight=W.cuda(), hessian_hook=hook, group_sizes=gs, group_bit_widths=bw, scale_bit_width=None,
                     gptq_use_kernel=use_kernel, gptq_block_sizes=gs, quant_symmetric=False, quant_mse=False)
    return hook, res


@pytest.mark.parametrize('name', GU.names())
def test_golden_end_to_end(name):
    from quant import reconstruct_nn_linear, W4A16Linear
    g = GU.load(name)
    hook, res = _run_gpu(g['W'], g['X'], g['bits'], g['act_order'], use_kernel=False)
    m = res['quant_meta']
    assert set(m) == {'qweight', 'scale', 'qzero', 'qscale', 'sscale', 'perm_inv', 'group_sizes', 'group_bit_widths',
                      'scale_bit_width'}
    assert m['qweight'].dtype == torch.uint8 and m['qzero'].dtype == torch.uint8
    assert m['scale'].dtype == g['dtype'] and m['perm_inv'].dtype == torch.int16
    assert m['group_sizes'].dtype == torch.int16 and m['group_bit_widths'].dtype == torch.uint8
    assert m['qweight'].device.type == 'cpu'            # save_device default
    assert torch.equal(m['perm_inv'], g['perm_inv'])
    match = O.code_match(m['qweight'], g['qweight'])
    assert match >= 0.999, match
    assert (m['qzero'].int() - g['qzero'].int()).abs().max() <= 1
    assert torch.allclose(m['scale'].float(), g['scale'], rtol=2e-3)
    for k in ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'):
        assert res['metrics'][k] == pytest.approx(g['metrics'][k], rel=2e-2), k
    # a8: dense reconstruction and the W4A16 kernel agree with the reference's W_hat
    lin = reconstruct_nn_linear(m, dtype=g['dtype'], device=torch.device('cuda'))
    x = torch.randn(5, g['C'], device='cuda').to(g['dtype'])
    y_ref = x.float() @ g['W_hat'].cuda().t()
    y_dense = lin(x).float()
    y_w4 = W4A16Linear.from_quant_meta(m, torch.device('cuda'))(x).float()
    for y in (y_dense, y_w4):
        rel = ((y - y_ref).norm() / y_ref.norm()).item()
        assert rel <= 1e-2, rel


@pytest.mark.parametrize('R,C,bits', [(512, 1024, 4), (1024, 2048, 4), (768, 1536, 3)])
def test_synthetic_vs_oracle(R, C, bits):
    W, X = O.synth_linear(R, C, 16, 512, seed=R + C + bits)
    ora = O.run_linear(W, X, 128, bits, 1e-2, True, mode='pure', batch=2)
    hook, res = _run_gpu(W, X, bits, True, use_kernel=False)
    assert O.rel_fro(hook.hessian.cpu(), ora['inv']['H']) <= 1e-3
    assert torch.equal(hook.perm.cpu(), ora['inv']['perm'])
    match = O.code_match(res['quant_meta']['qweight'], ora['result']['quant_meta']['qweight'])
    assert match >= 0.999, match
    for k in ('gptq_error', 'l2.0_error'):
        assert res['metrics'][k] == pytest.approx(ora['result']['metrics'][k], rel=1e-2), k
    # layer reconstruction error ||WX - W_hat X|| within 1% of the oracle's
    from quant import reconstruct_nn_linear
    Hp = ora['inv']['H']
    perm = ora['inv']['perm']
    q_gpu = reconstruct_nn_linear(res['quant_meta'], torch.float32).weight.data[:, perm]
    q_ora = reconstruct_nn_linear(ora['result']['quant_meta'], torch.float32).weight.data[:, perm]
    e_gpu = O.recon_error(W, q_gpu, Hp, perm)
    e_ora = O.recon_error(W, q_ora, Hp, perm)
    assert abs(e_gpu - e_ora) <= 0.01 * e_ora, (e_gpu, e_ora)


def test_kernel_path_semantics_vs_oracle():
    """gptq_use_kernel=True (what llama.py:156 passes): static scales + fp16 dequant."""
    W, X = O.synth_linear(256, 512, 8, 256, seed=5)
    ora = O.run_linear(W, X, 128, 4, 1e-2, True, mode='kernel', batch=2)
    hook, res = _run_gpu(W, X, 4, True, use_kernel=True)
    match = O.code_match(res['quant_meta']['qweight'], ora['result']['quant_meta']['qweight'])
    assert match >= 0.999, match
    assert torch.equal(res['quant_meta']['qzero'], ora['result']['quant_meta']['qzero'])


def test_unsupported_options_raise():
    from gptq_py import gptq_quant
    W = torch.zeros(8, 256, device='cuda', dtype=torch.float16)
    gs = torch.full([2], 128, dtype=torch.int32)
    bw = torch.full([2], 4, dtype=torch.int32)
    with pytest.raises(NotImplementedError):          # exllamav2's quantize_err (quant.py:131-150): absent from the reference too
        gptq_quant(W, None, gs, bw, quant_mse=True, quant_use_kernel=True)
    import types
    hook = types.SimpleNamespace(hessian_inv=torch.eye(256, device='cuda'), perm=None, perm_inv=None, hessian=None)
    with pytest.raises(NotImplementedError):          # sizes that are not multiples of 32 on the GPTQ branches
        gptq_quant(W, hook, gs, bw, gptq_use_kernel=False, quant_mse=False,
                   gptq_block_sizes=torch.tensor([100, 156], dtype=torch.int32))
    with pytest.raises(NotImplementedError):
        gptq_quant(W, hook, torch.tensor([72, 184], dtype=torch.int32), bw, gptq_use_kernel=True, quant_mse=False)
    # the round-to-nearest branch has no blocks (ref :101-124): gptq_block_sizes is ignored there, as in the reference
    a = gptq_quant(W, None, gs, bw, quant_mse=False, gptq_block_sizes=torch.full([4], 64, dtype=torch.int32), save_device=W.device)
    b = gptq_quant(W, None, gs, bw, quant_mse=False, gptq_block_sizes=gs, save_device=W.device)
    assert torch.equal(a['quant_meta']['qweight'], b['quant_meta']['qweight'])


def test_rtn_without_hook():
    """hessian_hook=None -> round to nearest (gptq_py.py:101-124)."""
    from gptq_py import gptq_quant
    W, _ = O.synth_linear(64, 256, 1, 8, seed=3)
    gs = torch.full([2], 128, dtype=torch.int32)
    bw = torch.full([2], 4, dtype=torch.int32)
    res = gptq_quant(W.cuda(), None, gs, bw, quant_mse=False)
    s, z, mq = O.find_params_asym(W.float().reshape(64, 2, 128), 4)
    want = O.quantize(W.float().reshape(64, 2, 128), s, z, mq).reshape(64, 256)
    assert torch.equal(res['quant_meta']['qweight'].float(), want)
    assert res['metrics']['gptq_error'] == 0.0


def test_deferred_metrics_equal_immediate_metrics():
    """gptq_quant(defer_metrics=True): same quant_meta, and a mapping whose values (read later, without a host sync inside
    the call) equal the floats of the default call; sharded.gptq_quant_group_sharded uses it for every linear of a group."""
    import gptq_py
    import sharded
    W, X = O.synth_linear(256, 512, 8, 256, seed=77)
    hook, res = _run_gpu(W, X, 4, True, use_kernel=False)
    gs = torch.full([4], 128, dtype=torch.int32)
    bw = torch.full([4], 4, dtype=torch.int32)
    kw = dict(group_sizes=gs, group_bit_widths=bw, scale_bit_width=None, gptq_use_kernel=False, gptq_block_sizes=gs,
              quant_symmetric=False, quant_mse=False, save_device=torch.device('cuda'))
    w = W.cuda()
    imm = gptq_py.gptq_quant(weight=w, hessian_hook=hook, **kw)
    lazy = gptq_py.gptq_quant(weight=w, hessian_hook=hook, defer_metrics=True, **kw)
    assert isinstance(lazy['metrics'], gptq_py.DeferredMetrics) and isinstance(imm['metrics'], dict)
    assert list(lazy['metrics']) == list(imm['metrics'])
    for k, v in imm['metrics'].items():
        assert lazy["metrics"][k] == pytest.approx(v, rel=1e-6), k      # colsq is an atomic fp32 sum: order varies run to run
    assert torch.equal(lazy['quant_meta']['qweight'], imm['quant_meta']['qweight'])
    grp = sharded.gptq_quant_group_sharded([w, w], hook, **kw)
    assert dict(grp[0]["metrics"]) == pytest.approx(imm["metrics"], rel=1e-6)
    assert torch.equal(grp[1]['quant_meta']['qweight'], imm['quant_meta']['qweight'])


@pytest.mark.parametrize('rows', [(256, 128, 384), (4096, 4096, 4096)])
def test_stacked_group_equals_separate_calls(rows):
    """gptq_quant_stacked (linears that share U solved as one row-stacked weight: what sharded.gptq_quant_group_sharded does
    for q / k / v on one rank) == one gptq_quant call per linear: identical codes / scales / zeros (rows are independent given
    U; at 3 x 4096 rows the stacked solve runs on 128-row tiles, the separate ones on 32-row tiles) and per-linear metrics
    equal to summation order."""
    import gptq_py
    import sharded
    C = 512
    W, X = O.synth_linear(sum(rows), C, 8, 256, seed=rows[0])
    hook, _ = _run_gpu(W[:64], X, 4, True, use_kernel=False)
    gs = torch.full([C // 128], 128, dtype=torch.int32)
    bw = torch.full([C // 128], 4, dtype=torch.int32)
    kw = dict(group_sizes=gs, group_bit_widths=bw, scale_bit_width=None, gptq_use_kernel=False, gptq_block_sizes=gs,
              quant_symmetric=False, quant_mse=False, save_device=torch.device('cuda'))
    ws, lo = [], 0
    for n in rows:
        ws.append(W[lo:lo + n].cuda().contiguous())
        lo += n
    sep = [gptq_py.gptq_quant(weight=w, hessian_hook=hook, **kw) for w in ws]
    stk = gptq_py.gptq_quant_stacked(ws, hook, **kw)
    assert sharded._stacking_pays([4096, 4096, 4096]) and not sharded._stacking_pays([11008, 11008])
    grp = sharded.gptq_quant_group_sharded(ws, hook, **kw)
    for a, b, c in zip(sep, stk, grp):
        for k in ('qweight', 'scale', 'qzero', 'perm_inv'):
            assert torch.equal(a['quant_meta'][k], b['quant_meta'][k]), k
            assert torch.equal(a['quant_meta'][k], c['quant_meta'][k]), k
        for k, v in a['metrics'].items():
            assert b['metrics'][k] == pytest.approx(v, rel=1e-5), k
            assert c['metrics'][k] == pytest.approx(v, rel=1e-5), k
