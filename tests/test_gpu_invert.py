"""K2 parity: fp32 GEMM building block, and damp + Cholesky + inverse factor vs the oracle."""
import pytest
import torch

import golden_util as GU
from oracle import gptq_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('flags', [0, 1, 1 | 2, 1 | 4, 8, 4])
@pytest.mark.parametrize('M,N,K', [(128, 128, 128), (256, 384, 128), (132, 260, 72), (512, 512, 512)])
def test_sgemm_flags(flags, M, N, K):
    import gptq
    torch.backends.cuda.matmul.allow_tf32 = False
    if flags & 2:
        N = M
    if flags & (4 | 8):
        M = N = K = max(M, K) // 4 * 4          # triangular operands are square here
    g = torch.Generator(device='cuda').manual_seed(flags * 1000 + M + N + K)
    ta = bool(flags & 1)
    A = torch.randn((K, M) if ta else (M, K), device='cuda', generator=g)
    B = torch.randn(K, N, device='cuda', generator=g)
    C0 = torch.randn(M, N, device='cuda', generator=g)
    Aeff, Beff = A, B
    if flags & 4:
        Aeff = A.triu()              # stored-form upper triangular; garbage below must be ignored
    if flags & 8:
        Beff = B.triu()
    opA = Aeff.t() if ta else Aeff
    want = 0.5 * C0 + (-1.5) * (opA.double() @ Beff.double()).float()
    C = C0.clone()
    gptq.sgemm(A, B, C, M, N, K, A.stride(0), B.stride(0), C.stride(0), -1.5, 0.5, flags)
    torch.cuda.synchronize()
    if flags & 2:
        # only tiles touching the upper triangle are defined
        assert torch.allclose(C.triu(), want.triu(), rtol=1e-4, atol=1e-3)
    else:
        assert torch.allclose(C, want, rtol=1e-4, atol=1e-3), (C - want).abs().max().item()


@pytest.mark.parametrize('D', [128, 192, 256, 640, 1024, 2304])
def test_invert_properties(D):
    """U upper, U^T U Hp = I, Hp = scale*H[perm][:,perm] + damp*I, on a synthetic SPD H."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(D)
    X = torch.randn(4 * D, D, device='cuda', generator=g) * torch.exp(torch.rand(D, device='cuda', generator=g) * 3 - 1.5)
    H = X.t() @ X
    perm = torch.randperm(D, device='cuda', generator=g)
    scale, damp = 2.0 / 16, 0.01 * float((H.diagonal() * 2.0 / 16).mean())
    Hup = H.triu()                       # K2 must only read the upper triangle
    Hp, U, info = gptq.hessian_invert(Hup, perm, scale, damp)
    torch.cuda.synchronize()
    assert int(info.item()) == 0
    want_hp = (H * scale)[perm][:, perm] + damp * torch.eye(D, device='cuda')
    assert torch.allclose(Hp, want_hp, rtol=1e-6, atol=0)
    assert torch.equal(U, U.triu())
    assert (U.diagonal() > 0).all()
    I = (U.double().t() @ U.double()) @ Hp.double()
    err = (I - torch.eye(D, device='cuda', dtype=torch.float64)).abs().max().item()
    assert err < 5e-3, err
    # against LAPACK in fp64
    L = torch.linalg.cholesky(want_hp.double())
    Uref = torch.linalg.cholesky(torch.cholesky_inverse(L), upper=True)
    rel = ((U.double() - Uref).norm() / Uref.norm()).item()
    assert rel < 1e-4, rel


def test_invert_flags_non_spd():
    import gptq
    D = 256
    H = -torch.eye(D, device='cuda')
    _, _, info = gptq.hessian_invert(H, None, 1.0, 0.0)
    assert int(info.item()) == 1


@pytest.mark.parametrize('name', GU.names())
def test_invert_golden(name):
    """HessianHook.invert vs U / perm / damped H frozen from the real reference."""
    from gptq_py import HessianHook
    g = GU.load(name)
    hook = HessianHook()
    hook.hessian = g['H_raw'].cuda()
    hook.n_samples = g['n_seq']
    U = hook.invert(1e-2, g['act_order'])
    if g['act_order']:
        assert torch.equal(hook.perm.cpu(), g['perm'])
        assert torch.equal(hook.perm_inv.cpu(), g['perm'].argsort())
    else:
        assert hook.perm is None
    assert torch.allclose(hook.hessian.cpu(), g['H_damped'], rtol=2e-6, atol=0)
    rel = O.rel_fro(U.cpu(), g['U'])
    assert rel < 1e-4, rel
    assert hook.invert_tries == 1
