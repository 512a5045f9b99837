"""Host-side mirror of the reference's quant.py (gptq-v2_b200/quant.py: construct_matrix, construct_matrix_2,
reconstruct_nn_linear -- plain torch, the a8 path of SURVEY section 8) pinned on CPU against the dense reconstructions
the REAL reference produced for the 12 golden cases (tests/golden/, oracle/gen_golden.py)."""
import pytest
import torch

import golden_util as GU


def _meta(g, G):
    dt = g['dtype']
    m = {'qweight': g['qweight'].to(torch.uint8), 'scale': g['scale'].to(dt), 'qzero': g['qzero'].to(torch.uint8),
         'perm_inv': g['perm_inv'], 'group_sizes': torch.full([G], 128, dtype=torch.int16),
         'qscale': None, 'sscale': None, 'scale_bit_width': None}
    if g.get('qscale') is not None:
        m['qscale'], m['sscale'] = g['qscale'].to(torch.uint8), g['sscale'].to(dt)
    return m


@pytest.mark.parametrize('name', GU.names() + GU.qnames())
def test_reconstruct_nn_linear_matches_reference(name):
    import quant
    g = GU.load(name) if name.startswith('r') else GU.loadq(name)
    G = g['C'] // 128
    lin = quant.reconstruct_nn_linear(_meta(g, G), g['dtype'], torch.device('cpu'))
    assert lin.weight.dtype == g['dtype'] and tuple(lin.weight.shape) == (g['R'], g['C'])
    assert torch.equal(lin.weight.data.float(), g['W_hat'])
    # and nn.Linear semantics: y = x W_hat^T, no bias
    x = torch.randn(3, g['C'], generator=torch.Generator().manual_seed(1)).to(g['dtype'])
    assert lin.bias is None
    assert torch.equal(lin(x), x @ lin.weight.data.t())
