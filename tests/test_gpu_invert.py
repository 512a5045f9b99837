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


@pytest.mark.parametrize('M,N,K,upper', [(128, 256, 16, False), (128, 256, 128, False), (256, 512, 128, False),
                                         (3968, 4096, 128, False), (200, 300, 64, False), (1024, 1024, 128, True),
                                         (640, 640, 64, True)])
def test_gemm_tf32x3_is_fp32_faithful(M, N, K, upper):
    """C += alpha A^T B on tcgen05 with the 3xTF32 split: error must be at fp32-reordering level,
    far below a single TF32 pass (~5e-4)."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(M + N + K)
    lda, ldb, ldc = M + 64, N + 32, N + 16                     # operands are windows of larger matrices
    A = torch.randn(K, lda, device='cuda', generator=g)
    B = torch.randn(K, ldb, device='cuda', generator=g)
    C0 = torch.randn(M, ldc, device='cuda', generator=g)
    C = C0.clone()
    gptq.gemm_tf32x3(A, B, C, M, N, K, lda, ldb, ldc, -1.0, upper)
    torch.cuda.synchronize()
    want = C0[:, :N].double() - A[:, :M].double().t() @ B[:, :N].double()
    got = C[:, :N].double()
    assert torch.equal(C[:, N:], C0[:, N:])                  # nothing outside the window is touched
    if upper:
        tm, tn = 128, 256
        mask = torch.zeros(M, N, dtype=torch.bool, device='cuda')
        for m0 in range(0, M, tm):
            for n0 in range(0, N, tn):
                if not (m0 >= n0 + tn):
                    mask[m0:m0 + tm, n0:n0 + tn] = True
        assert torch.equal(C[:, :N][~mask], C0[:, :N][~mask])
        err = ((got - want)[mask]).abs().max().item()
    else:
        err = (got - want).abs().max().item()
    scale = (A[:, :M].abs().double().t() @ B[:, :N].abs().double()).max().item()
    assert err <= 4e-6 * scale, (err, scale)


@pytest.mark.parametrize('M,N,K,upper', [(3968, 4096, 128, False), (2200, 2560, 192, False), (4096, 4096, 1024, False),
                                         (10880, 10880, 128, True), (2304, 2304, 64, True)])
def test_gemm_tf32x3_cta_pair(M, N, K, upper):
    """Mode 3: 256 x 256 tiles on CTA pairs (tcgen05 cta_group::2, each CTA stages half of B) whenever there is more than
    one wave of tiles.  Same fp32-faithful bound; ragged M (the second CTA of the last pair partly or wholly past the
    edge), upper-only tile skipping on 256-row tiles."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(M + N + K + 3)
    A = torch.randn(K, M, device='cuda', generator=g)
    B = torch.randn(K, N, device='cuda', generator=g)
    C0 = torch.randn(M, N, device='cuda', generator=g)
    prev = gptq.set_tf32x3(3)
    try:
        C = C0.clone()
        gptq.gemm_tf32x3(A, B, C, M, N, K, M, N, N, -1.0, upper)
        torch.cuda.synchronize()
    finally:
        gptq.set_tf32x3(prev)
    want = C0.double() - A.double().t() @ B.double()
    scale = (A.abs().double().t() @ B.abs().double()).max().item()
    if upper:
        mask = torch.zeros(M, N, dtype=torch.bool, device='cuda')
        for m0 in range(0, M, 256):
            for n0 in range(0, N, 256):
                if not (m0 >= n0 + 256):
                    mask[m0:m0 + 256, n0:n0 + 256] = True
        assert torch.equal(C[~mask], C0[~mask])
        triu = torch.ones(M, N, dtype=torch.bool, device='cuda').triu()
        assert (mask | ~triu).all()                           # everything on or above the diagonal is computed
        err = ((C.double() - want)[mask]).abs().max().item()
    else:
        err = (C.double() - want).abs().max().item()
    assert err <= 4e-6 * scale, (err, scale)


@pytest.mark.parametrize('D', [4096, 11008])
def test_invert_cta_pair_mode_matches(D):
    """hessian_invert with the CTA-pair GEMM (far updates of the Cholesky and, from D = 11008 on, the batched triangular
    products with their clipped k ranges) == the single-CTA GEMM up to fp32 summation order, and U^T U H = I."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(5)
    X = torch.randn(2 * D, D, device='cuda', generator=g) * torch.exp(torch.rand(D, device='cuda', generator=g) * 3 - 1.5)
    H = X.t() @ X
    perm = torch.argsort(H.diagonal(), descending=True)
    damp = 0.01 * float(H.diagonal().mean())
    outs = {}
    for mode in (2, 3):
        prev = gptq.set_tf32x3(mode)
        try:
            _, U, info = gptq.hessian_invert(H, perm, 1.0, damp)
            torch.cuda.synchronize()
        finally:
            gptq.set_tf32x3(prev)
        assert int(info.item()) == 0
        outs[mode] = U
    rel = ((outs[3].double() - outs[2].double()).norm() / outs[2].double().norm()).item()
    assert rel <= 1e-5, rel
    # independent check of the pair-mode result on a slice: rows of (U^T U) (H_perm + damp I) = I
    Hp = (H[perm][:, perm] + damp * torch.eye(D, device='cuda')).double()
    Ud = outs[3].double()
    rows = torch.arange(0, D, max(1, D // 64), device='cuda')
    got = (Ud.t()[rows] @ Ud) @ Hp
    want = torch.zeros_like(got)
    want[torch.arange(len(rows), device='cuda'), rows] = 1.0
    assert (got - want).abs().max().item() <= 2e-2


@pytest.mark.parametrize('M,N,K', [(128, 256, 128), (3968, 4096, 128), (1024, 768, 1024)])
def test_gemm_tf32x3_truncation_split(M, N, K):
    """Mode 2: the landed fp32 tile is the hi operand as it is (the tensor core reads its upper 19 bits) and only
    lo = rna_tf32(x - trunc_tf32(x)) is written.  Same fp32-faithful bound as the round-to-nearest split."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(M + N + K + 1)
    A = torch.randn(K, M, device='cuda', generator=g)
    B = torch.randn(K, N, device='cuda', generator=g)
    C0 = torch.randn(M, N, device='cuda', generator=g)
    want = C0.double() - A.double().t() @ B.double()
    scale = (A.abs().double().t() @ B.abs().double()).max().item()
    errs = {}
    for mode in (1, 2):
        prev = gptq.set_tf32x3(mode)
        try:
            C = C0.clone()
            gptq.gemm_tf32x3(A, B, C, M, N, K, M, N, N, -1.0, False)
            torch.cuda.synchronize()
        finally:
            gptq.set_tf32x3(prev)
        errs[mode] = (C.double() - want).abs().max().item()
    assert errs[1] <= 4e-6 * scale and errs[2] <= 4e-6 * scale, (errs, scale)


def test_tf32x3_switch_gives_same_codes():
    """K4 on tensor cores (3xTF32) vs fp32 SIMT: the quantisation codes must agree to >= 99.9%."""
    import gptq
    W, X = O.synth_linear(512, 1024, 8, 512, seed=77)
    H, n = O.hessian_accumulate(None, 0, X)
    inv = O.hessian_invert(H, n, 1e-2, True)
    U = inv['U'].cuda().contiguous()
    outs = []
    for enable in (1, 0, 2):
        prev = gptq.set_tf32x3(enable)
        try:
            Wt = gptq.prep_weight_t(W.cuda().contiguous(), inv['perm'].cuda())
            _, codes_t, _, _, _ = gptq.gptq_solve(Wt, U, 128, 15.0)
            outs.append(codes_t.cpu())
        finally:
            gptq.set_tf32x3(prev)
    assert (outs[0] == outs[1]).float().mean().item() >= 0.999
    assert (outs[2] == outs[1]).float().mean().item() >= 0.999


@pytest.mark.parametrize('D', [128, 200, 1000, 4096])
def test_pack_unpack_upper_tiles(D):
    """The multi-GPU reduction payload: upper 128 x 128 tiles of H packed into one contiguous buffer and back, bit for
    bit; everything outside those tiles is left alone; a sum over two 'ranks' done on the packed buffers equals the sum
    done on the matrices wherever invert() reads (the upper triangle)."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(D)
    H1 = torch.randn(D, D, device='cuda', generator=g)
    H2 = torch.randn(D, D, device='cuda', generator=g)
    nb = (D + 127) // 128
    p1, p2 = gptq.pack_upper(H1), gptq.pack_upper(H2)
    assert p1.numel() == nb * (nb + 1) // 2 * 128 * 128
    out = torch.full((D, D), -7.0, device='cuda')
    gptq.unpack_upper(p1, out)
    torch.cuda.synchronize()
    tile_mask = torch.zeros(D, D, dtype=torch.bool, device='cuda')
    for bi in range(nb):
        tile_mask[bi * 128:(bi + 1) * 128, bi * 128:] = True
    assert torch.equal(out[tile_mask], H1[tile_mask])
    assert (out[~tile_mask] == -7.0).all()
    Hs = H1.clone()
    gptq.unpack_upper(p1 + p2, Hs)
    triu = torch.ones(D, D, dtype=torch.bool, device='cuda').triu()
    assert torch.equal(Hs[triu], (H1 + H2)[triu])


@pytest.mark.parametrize('D', [256, 1408])
def test_invert_damping_from_device_scalar(D):
    """gq_hessian_invert_dev (damping read on the device, what HessianHook.invert uses) == gq_hessian_invert with the same
    value passed from the host, bit for bit, for the first try and for a cumulative retry (damp * 3)."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(D + 5)
    X = torch.randn(3 * D, D, device='cuda', generator=g)
    H = (X.t() @ X).triu()
    perm = torch.randperm(D, device='cuda', generator=g)
    damp_dev = (0.01 * H.diagonal().mean()).reshape(())
    for mult in (1.0, 3.0):
        Hp_a, U_a, info_a = gptq.hessian_invert(H, perm, 0.25, damp_dev, damp_mult=mult)
        Hp_b, U_b, info_b = gptq.hessian_invert(H, perm, 0.25, float((damp_dev * mult).item()))
        torch.cuda.synchronize()
        assert int(info_a.item()) == 0 and int(info_b.item()) == 0
        assert torch.equal(Hp_a, Hp_b) and torch.equal(U_a, U_b)
    with pytest.raises(gptq.GptqB200Error):
        gptq.hessian_invert(H, perm, 0.25, damp_dev.double())


def test_programmatic_dependent_launch_changes_nothing():
    """gq_set_pdl(0 / 1): the same kernels in the same order, only their set-up overlaps the predecessor's tail; U, Hp and the
    solver outputs are bit-identical."""
    import gptq
    D = 1408
    g = torch.Generator(device='cuda').manual_seed(3)
    X = torch.randn(3 * D, D, device='cuda', generator=g)
    H = (X.t() @ X).triu()
    perm = torch.randperm(D, device='cuda', generator=g)
    Wt0 = torch.randn(D, 192, device='cuda', generator=g) * 0.02
    outs = []
    for pdl in (1, 0):
        old = gptq.set_pdl(pdl)
        try:
            Hp, U, info = gptq.hessian_invert(H, perm, 0.5, 0.01 * float(H.diagonal().mean()))
            sol = gptq.gptq_solve(Wt0.clone(), U, 128, 15.0)
            torch.cuda.synchronize()
        finally:
            gptq.set_pdl(old)
        assert int(info.item()) == 0
        outs.append((Hp, U) + tuple(sol))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
