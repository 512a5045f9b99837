"""K1 parity: tcgen05 SYRK vs fp32 X^T X (the reference's use_kernel=False path, gptq_py.py:31-33)."""
import pytest
import torch

import golden_util as GU

pytestmark = pytest.mark.gpu


def _ref(H0, X):
    torch.backends.cuda.matmul.allow_tf32 = False
    x = X.float()
    return H0 + x.t() @ x


# tcgen05 accumulates in fp32 with truncation, like every tensor-core generation: the error
# grows ~linearly with the number of accumulation steps (T/16).  Requirement (BASELINE.json): 1e-3.
TOL = 3e-4


def _relfro(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


@pytest.mark.parametrize('variant', [1, 2])
# (6400, 192) and (11008, 128): 25 / 43 super-tiles per side, the blocked (8 x 8) tile walk of decode_tile
@pytest.mark.parametrize('D,T', [(256, 64), (256, 200), (512, 1024), (64, 1), (320, 100), (1024, 4096), (4096, 2048),
                                 (6400, 192), (11008, 128)])
@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
def test_syrk_matches_fp32(variant, D, T, dtype):
    import gptq
    g = torch.Generator(device='cuda').manual_seed(D * 7 + T)
    X = torch.randn(T, D, device='cuda', generator=g).to(dtype)
    S = torch.randn(D, D, device='cuda', generator=g)
    H0 = S + S.t()
    want = _ref(H0, X)
    # mode 2: both triangles
    H = H0.clone()
    gptq.hessian_syrk(H, X, mode=2, variant=variant)
    torch.cuda.synchronize()
    err = _relfro(H, want)
    assert err <= 1e-3, f'mode2 variant{variant} D{D} T{T}: rel fro {err:.3e}; max abs {(H - want).abs().max().item():.3e}'
    assert err <= TOL
    # mode 1: upper + mirror -> exactly symmetric
    H = H0.clone()
    gptq.hessian_syrk(H, X, mode=1, variant=variant)
    assert torch.equal(H, H.t())
    assert _relfro(H, want) <= TOL
    # mode 0: strictly-lower part untouched, upper correct
    H = H0.clone()
    gptq.hessian_syrk(H, X, mode=0, variant=variant)
    assert torch.equal(H.tril(-1), H0.tril(-1))
    assert _relfro(H.triu(), want.triu()) <= TOL


def test_syrk_asymmetric_h_like_reference_test():
    """package/tests/test.py: random (asymmetric) H, 256 x 512 fp16 input, h + x32^T x32."""
    import gptq
    x = torch.randn(256, 512, dtype=torch.float16, device='cuda')
    h = torch.randn(512, 512, dtype=torch.float32, device='cuda')
    want = _ref(h, x)
    gptq.accumulate_hessian(h, x)
    assert (h - want).abs().max().item() < 1e-3


def test_syrk_rejects_other_dtypes():
    import gptq
    h = torch.zeros(64, 64, device='cuda')
    with pytest.raises(NotImplementedError):
        gptq.accumulate_hessian(h, torch.zeros(8, 64, device='cuda'))


@pytest.mark.parametrize('name', GU.names())
def test_syrk_golden_hessian(name):
    """H accumulated over the fixture's batches vs the H frozen from the real reference."""
    from gptq_py import HessianHook
    g = GU.load(name)
    hook = HessianHook()
    X = g['X'].cuda()
    for i in range(0, g['n_seq'], 2):
        hook.add_batch(X[i:i + 2])
    assert hook.n_samples == g['n_seq']
    # HessianHook accumulates the upper triangle only (invert() reads nothing else)
    err = _relfro(hook.hessian.cpu().triu(), g['H_raw'].triu())
    assert err <= TOL, err


def test_triton_unit_test_properties():
    """gptq_triton/accumulate_hessian.py:161-192: T=16384, D=4096, seed 0, symmetric H0;
    mirrored result exactly symmetric and equal on the lower triangle."""
    from gptq_triton.accumulate_hessian import accumulate_hessian
    torch.manual_seed(0)
    X = torch.randn(16384, 4096, device='cuda', dtype=torch.float16)
    S = torch.randn(4096, 4096, device='cuda')
    H0 = S + S.t()
    a, b = H0.clone(), H0.clone()
    accumulate_hessian(a, X, save_lower_only=True)
    accumulate_hessian(b, X, save_lower_only=False)
    assert torch.equal(a.tril(), b.tril())
    assert torch.equal(b, b.t())
    want = _ref(H0, X)
    assert _relfro(b, want) <= TOL
