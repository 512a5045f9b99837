"""gptq-v2_b200/quant.py's Quantizer (the host mirror of the reference class, quant.py:72-183) against the oracle's
restatement (oracle/quantizer_oracle.py, itself pinned bit-exact on reference-generated golden cases): same scales, zero
points, EXL2 scale grids and MSE choices, bit for bit, on CPU."""
import pytest
import torch

from oracle import gptq_oracle as G
from oracle import quantizer_oracle as Q


@pytest.mark.parametrize('bits', [3, 4])
@pytest.mark.parametrize('sym,scale_bits,mse', [(False, None, False), (True, None, False), (False, None, True),
                                                (True, None, True), (False, 4, False), (False, 4, True)])
def test_quantizer_host_mirror_matches_oracle(bits, sym, scale_bits, mse):
    import quant
    g = torch.Generator().manual_seed(bits * 100 + int(sym) * 10 + int(mse) + (scale_bits or 0))
    x = torch.randn(48, 128, generator=g) * torch.exp(torch.rand(48, 1, generator=g) * 2 - 1) * 0.02
    P = Q.find_params(x, bits, sym, scale_bits)
    if mse:
        P = Q.mse_refine(P, x, 0.8, 100, 2.4)
    qz = quant.Quantizer()
    qz.find_params(x, bit_width=bits, sym=sym, scale_bit_width=scale_bits)
    if mse:
        qz.mse(x, max_shrink=0.8, n_grid=100, norm=2.4, use_kernel=False)
    assert torch.equal(qz.scale, P['scale']) and torch.equal(qz.qzero, P['qzero'])
    assert float(qz.maxq) == float(P['maxq'])
    if scale_bits is not None:
        assert torch.equal(qz.qscale, P['qscale']) and torch.equal(qz.sscale.reshape(-1), P['sscale'].reshape(-1))
    codes = qz.quantize(x)
    want = G.quantize(x, P['scale'], P['qzero'], P['maxq'])
    assert torch.equal(codes, want)
    assert torch.equal(qz.dequantize(codes), G.dequantize(want, P['scale'], P['qzero']))


def test_quantizer_mse_kernel_option_is_refused():
    import quant
    qz = quant.Quantizer()
    x = torch.randn(4, 128)
    qz.find_params(x, bit_width=4)
    with pytest.raises(NotImplementedError):
        qz.mse(x, use_kernel=True)
