"""K5 bit-exact pack/unpack and K6 W4A16 GEMM vs the dense reconstruction (quant.py:262-279)."""
import numpy as np
import pytest
import torch

from oracle import gptq_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('R,C', [(1, 8), (37, 256), (128, 4096), (1000, 11008)])
def test_pack_unpack_bit_exact(R, C):
    import gptq
    rng = np.random.default_rng(R * C)
    codes = rng.integers(0, 16, size=(R, C), dtype=np.uint8)
    want = O.pack_int4(codes)
    got = gptq.pack_int4(torch.from_numpy(codes).cuda())
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want)
    back = gptq.unpack_int4(got)
    assert np.array_equal(back.cpu().numpy(), codes)
    c2, p2 = gptq.transpose_pack_int4(torch.from_numpy(np.ascontiguousarray(codes.T)).cuda())
    assert np.array_equal(c2.cpu().numpy(), codes)
    assert np.array_equal(p2.cpu().numpy().view(np.uint32), want)


def test_pack_empty():
    import gptq
    out = gptq.pack_int4(torch.empty(0, 64, dtype=torch.uint8, device='cuda'))
    assert out.shape == (0, 8)


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M', [1, 2, 3, 4, 7, 16])
@pytest.mark.parametrize('N,K', [(256, 512), (4096, 4096), (1024, 11008)])
def test_w4a16_vs_dense(M, N, K, dtype):
    """rel err <= 1e-2 vs x @ reconstruct_dense(meta)^T (the reference's evaluation path)."""
    import gptq
    g = torch.Generator().manual_seed(M + N + K)
    G = K // 128
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    perm = torch.randperm(K, generator=g)
    x = torch.randn(M, K, generator=g).to(dtype)
    meta = dict(qweight=codes, scale=scale, qzero=qzero, perm_inv=perm.argsort().to(torch.int16) if K < 32768 else None,
                group_sizes=torch.full((G,), 128, dtype=torch.int16))
    want = O.linear_forward_dense(x, meta, dtype)
    packed = gptq.pack_int4(codes.cuda())
    y = gptq.w4a16_gemm(x.cuda(), packed, scale.t().contiguous().cuda(), qzero.t().contiguous().cuda(),
                        perm.to(torch.int32).cuda(), 128)
    torch.cuda.synchronize()
    rel = ((y.float().cpu() - want).norm() / want.norm()).item()
    assert rel <= 1e-2, rel
    want2 = O.w4a16_reference(x, O.pack_int4(codes.numpy()), scale, qzero, perm, 128)
    rel2 = ((y.float().cpu() - want2).norm() / want2.norm()).item()
    assert rel2 <= 5e-3, rel2


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M,N,K,group', [(17, 200, 384, 128), (33, 1000, 1024, 64), (64, 640, 2048, 32), (65, 512, 768, 128),
                                         (130, 384, 4096, 128), (300, 136, 288, 32)])
def test_w4a16_tcgen05_ragged_shapes(M, N, K, group, dtype):
    """batch tiles of 64 with padding, weight rows not a multiple of 128, K not a multiple
    of 256, groups of 32/64/128, no permutation."""
    import gptq
    g = torch.Generator().manual_seed(M * 7 + N + K)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    x = torch.randn(M, K, generator=g).to(dtype)
    meta = dict(qweight=codes, scale=scale, qzero=qzero, perm_inv=None, group_sizes=torch.full((G,), group, dtype=torch.int16))
    want = O.linear_forward_dense(x, meta, dtype)
    y = gptq.w4a16_gemm(x.cuda(), gptq.pack_int4(codes.cuda()), scale.t().contiguous().cuda(),
                        qzero.t().contiguous().cuda(), None, group)
    torch.cuda.synchronize()
    rel = ((y.float().cpu() - want).norm() / want.norm()).item()
    assert rel <= 1e-2, rel


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M,N,K,group,use_perm', [(1, 592, 384, 128, True), (2, 4096, 4096, 128, True), (3, 16, 128, 128, False),
                                                  (5, 1024, 11008, 128, True), (8, 2048, 2048, 256, False),
                                                  (1, 28672, 8192, 128, True), (8, 320, 1024, 512, True)])
def test_w4a16_small_batch_kernel(M, N, K, group, use_perm, dtype):
    """Batch 1..8 path (w4a16_smallm.cu): offset-carrying mma fragments + per-chunk correction through the activation
    sums.  Activations with a large mean make that cancellation visible; the bound is the 16-bit rounding of y."""
    import gptq
    g = torch.Generator().manual_seed(M * 131 + N + K + group)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    perm = torch.randperm(K, generator=g) if use_perm else None
    x = (torch.randn(M, K, generator=g) + 1.5).to(dtype)
    # float64 ground truth on the same 16-bit inputs: y = sum_k (q - z) s x[perm[k]]
    w = (codes.double() - qzero.double().repeat_interleave(group, dim=1)) * scale.double().repeat_interleave(group, dim=1)
    xp = x.double()[:, perm] if use_perm else x.double()
    want = xp @ w.t()
    y = gptq.w4a16_gemm(x.cuda(), gptq.pack_int4(codes.cuda()), scale.t().contiguous().cuda(), qzero.t().contiguous().cuda(),
                        perm.to(torch.int32).cuda() if use_perm else None, group)
    torch.cuda.synchronize()
    rel = ((y.double().cpu() - want).norm() / want.norm()).item()
    assert rel <= (6e-4 if dtype == torch.float16 else 4e-3), rel
    # the fp32 sums themselves (before the 16-bit store) must be far tighter: re-round the truth and compare elementwise
    # elementwise: one 16-bit ulp of the result plus 2e-6 of the absolute sum (fp32 accumulation incl. the offset term)
    bound = want.abs() * (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7) + 2e-6 * (xp.abs() @ w.abs().t())
    assert ((y.double().cpu() - want).abs() <= bound).all()


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M,N,K,group,use_perm', [(9, 256, 256, 128, True), (16, 4096, 4096, 128, True), (12, 1008, 11008, 128, True),
                                                  (16, 2048, 2048, 256, False), (11, 592, 384, 128, True), (13, 320, 1024, 512, True),
                                                  (4, 2048, 28672, 128, True), (8, 64, 16512, 128, False), (16, 28672, 8192, 128, True),
                                                  (10, 16, 128, 128, False), (1, 2064, 4224, 128, True)])
def test_w4a16_mid_batch_kernel(M, N, K, group, use_perm, dtype):
    """Batch 9..16 (and batch 3..8 on rows too long for the small-batch kernel) path (w4a16_midm.cu): the small-batch
    arithmetic with 1 / 2 blocks of 8 batch columns and the activation streamed through shared memory in windows (full and
    ragged last windows, one to many rounds, CTAs with 1 .. 13 row tiles, warps without units in a round).  Forced through
    set_w4a16_path(4) so that the shapes the automatic dispatch keeps on other kernels run it too.  Same bounds as the
    small-batch test; the tcgen05 route (set_w4a16_path(3): the dispatch without this kernel) must agree to the 16-bit
    rounding."""
    import gptq
    g = torch.Generator().manual_seed(M * 137 + N + K + group)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    perm = torch.randperm(K, generator=g) if use_perm else None
    x = (torch.randn(M, K, generator=g) + 1.5).to(dtype)
    w = (codes.double() - qzero.double().repeat_interleave(group, dim=1)) * scale.double().repeat_interleave(group, dim=1)
    xp = x.double()[:, perm] if use_perm else x.double()
    want = xp @ w.t()
    args = (x.cuda(), gptq.pack_int4(codes.cuda()), scale.t().contiguous().cuda(), qzero.t().contiguous().cuda(),
            perm.to(torch.int32).cuda() if use_perm else None, group)
    old = gptq.set_w4a16_path(4)
    try:
        y = gptq.w4a16_gemm(*args)
        y2 = gptq.w4a16_gemm(*args)                # a second call: equal up to the order of the shared-memory adds of a tile's
        torch.cuda.synchronize()                   # partial sums
        gptq.set_w4a16_path(0)
        y_auto = gptq.w4a16_gemm(*args)            # what the automatic dispatch picks for this shape
        gptq.set_w4a16_path(3)
        y_tc = gptq.w4a16_gemm(*args)
        torch.cuda.synchronize()
    finally:
        gptq.set_w4a16_path(old)
    rel = ((y.double().cpu() - want).norm() / want.norm()).item()
    assert rel <= (6e-4 if dtype == torch.float16 else 4e-3), rel
    bound = want.abs() * (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7) + 2e-6 * (xp.abs() @ w.abs().t())
    assert ((y.double().cpu() - want).abs() <= bound).all()
    assert ((y2.double().cpu() - want).abs() <= bound).all()
    for other in (y_auto, y_tc):
        rel_o = ((y.double() - other.double()).norm() / other.double().norm()).item()
        assert rel_o <= (1.2e-3 if dtype == torch.float16 else 8e-3), rel_o


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M,N,K,group,use_perm', [(9, 256, 256, 128, True), (16, 200, 320, 64, False), (33, 1000, 1088, 32, True),
                                                  (64, 4096, 4096, 128, True), (100, 520, 2048, 128, False),
                                                  (128, 1024, 11008, 128, True), (24, 28672, 512, 128, False)])
def test_w4a16_weights_on_n_kernel(M, N, K, group, use_perm, dtype):
    """Weights-on-N tcgen05 kernel (w4a16_wb.cu, opt-in through set_w4a16_path(2)): weight rows not a multiple of 256, K not a multiple of 256, k-split units,
    zero-filled batch padding; vs the dense reconstruction (a8)."""
    import gptq
    g = torch.Generator().manual_seed(M * 17 + N + K + group)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    perm = torch.randperm(K, generator=g) if use_perm else None
    x = torch.randn(M, K, generator=g).to(dtype)
    meta = dict(qweight=codes, scale=scale, qzero=qzero, perm_inv=perm.argsort().to(torch.int16) if use_perm else None,
                group_sizes=torch.full((G,), group, dtype=torch.int16))
    want = O.linear_forward_dense(x, meta, dtype)
    old = gptq.set_w4a16_path(2)
    try:
        y = gptq.w4a16_gemm(x.cuda(), gptq.pack_int4(codes.cuda()), scale.t().contiguous().cuda(), qzero.t().contiguous().cuda(),
                            perm.to(torch.int32).cuda() if use_perm else None, group)
        torch.cuda.synchronize()
    finally:
        gptq.set_w4a16_path(old)
    rel = ((y.float().cpu() - want).norm() / want.norm()).item()
    assert rel <= 1e-2, rel


@pytest.mark.parametrize('M', [1, 5, 8])
def test_w4a16_paths_agree(M):
    """The same small-batch problem through the warp-level mma kernel (default) and through the tcgen05 kernels."""
    import gptq
    g = torch.Generator().manual_seed(M)
    N, K, group = 2048, 4096, 128
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(K // group, N, generator=g) * 0.01 + 1e-3).half().cuda()
    qzero = torch.randint(0, 16, (K // group, N), generator=g, dtype=torch.uint8).cuda()
    x = torch.randn(M, K, generator=g).half().cuda()
    packed = gptq.pack_int4(codes.cuda())
    ys = []
    for path in (0, 2, 1):
        old = gptq.set_w4a16_path(path)
        try:
            ys.append(gptq.w4a16_gemm(x, packed, scale, qzero, None, group).float())
        finally:
            gptq.set_w4a16_path(old)
    torch.cuda.synchronize()
    for y in ys[1:]:
        assert ((y - ys[0]).norm() / ys[0].norm()).item() <= 5e-3


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize('R,C', [(64, 128), (100, 200), (4096, 4096), (36, 1030)])
def test_prep_weight_scatter_form_matches_gather_form(dtype, R, C):
    """gq_prep_weight_t3 (inverse permutation, coalesced scatter) == gq_prep_weight_t2 (permutation, gather): Wt and
    orig_t bit for bit, colsq to fp32 summation order; and both equal W[:, perm].t()."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(R * 7 + C)
    W = (torch.randn(R, C, device='cuda', generator=g) * 0.02).to(dtype)
    perm = torch.randperm(C, device='cuda', generator=g)
    pinv = torch.argsort(perm)
    Wt_g, o_g, sq_g = gptq.prep_weight_t2(W, perm)
    Wt_s, o_s, sq_s = gptq.prep_weight_t2(W, perm, pinv)
    torch.cuda.synchronize()
    want = W[:, perm].t().contiguous()
    assert torch.equal(Wt_s, want.float()) and torch.equal(Wt_g, Wt_s)
    assert torch.equal(o_s, want) and torch.equal(o_g, o_s)
    ref_sq = (want.double() ** 2).sum(dim=1)
    assert torch.allclose(sq_s.double(), ref_sq, rtol=1e-5, atol=1e-12)
    assert torch.allclose(sq_g.double(), ref_sq, rtol=1e-5, atol=1e-12)
    # no permutation: identity scatter
    Wt_i, o_i, sq_i = gptq.prep_weight_t2(W, None)
    assert torch.equal(Wt_i, W.t().float().contiguous()) and torch.equal(o_i, W.t().contiguous())


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('N,K,group', [(64, 128, 128), (100, 384, 64), (4096, 4096, 128), (24, 1024, 32)])
def test_w4a16_dequant_matches_reference_construct_matrix(N, K, group, dtype):
    """gq_w4a16_dequant == the reference's construct_matrix (quant.py:222-237: int16 subtract, multiply in the 16-bit
    dtype) bit for bit, in the permuted column order of quant_meta."""
    import gptq
    import quant
    g = torch.Generator().manual_seed(N + K + group)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale = (torch.rand(N, G, generator=g) * 0.01 + 1e-3).to(dtype)
    qzero = torch.randint(0, 16, (N, G), generator=g, dtype=torch.uint8)
    want = quant.construct_matrix(codes, scale, qzero, torch.full([G], group, dtype=torch.int16))
    got = gptq.w4a16_dequant(gptq.pack_int4(codes.cuda()), scale.t().contiguous().cuda(), qzero.t().contiguous().cuda(), group)
    torch.cuda.synchronize()
    assert got.dtype == dtype and torch.equal(got.cpu(), want)


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('R,C,group,use_perm', [(64, 128, 128, True), (100, 384, 64, False), (4096, 4096, 128, True),
                                                (24, 1024, 32, True), (130, 11008, 128, True), (16, 240, 48, True)])
def test_reconstruct_nn_linear_fused_route_matches_torch_route(R, C, group, use_perm, dtype):
    """reconstruct_nn_linear on CUDA with the group-size hint (gq_reconstruct_dense: construct_matrix + the perm_inv gather
    in one launch) == the same function's torch route (device-resident group_sizes, no hint) == the oracle's restatement
    of quant.py:222-237, 262-279, bit for bit; any group size that divides C, with and without act-order."""
    import quant
    g = torch.Generator().manual_seed(R + C + group)
    G = C // group
    meta = dict(qweight=torch.randint(0, 16, (R, C), generator=g, dtype=torch.uint8),
                scale=(torch.rand(R, G, generator=g) * 0.01 + 1e-3).to(dtype),
                qzero=torch.randint(0, 16, (R, G), generator=g, dtype=torch.uint8),
                perm_inv=torch.randperm(C, generator=g).to(torch.int16) if use_perm else None,
                group_sizes=torch.full((G,), group, dtype=torch.int16),
                group_bit_widths=torch.full((G,), 4, dtype=torch.int16))
    want = O.reconstruct_dense(meta, dtype)
    dev = torch.device('cuda')
    on_gpu = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in meta.items()}
    fused = quant.reconstruct_nn_linear(on_gpu, dtype, dev, group_size=group)
    assert quant._fused_reconstruction(on_gpu, dtype, dev, group) is not None          # the hint takes the kernel
    assert quant._fused_reconstruction(on_gpu, dtype, dev, None) is None               # device group_sizes, no hint: torch route
    assert quant._fused_reconstruction(meta, dtype, dev, None) is not None             # host group_sizes: read for free
    plain = quant.reconstruct_nn_linear(on_gpu, dtype, dev)
    torch.cuda.synchronize()
    assert fused.weight.dtype == dtype and fused.weight.shape == (R, C) and fused.in_features == C and fused.out_features == R
    assert torch.equal(fused.weight.data.cpu(), want)
    assert torch.equal(plain.weight.data.cpu(), want)


@pytest.mark.parametrize('M', [128, 600])
def test_w4a16_linear_large_batch_route(M):
    """W4A16Linear switches to one dequantisation + a dense GEMM for compute-bound batches; both routes agree with the
    dense reconstruction of the reference (a8) within the 1e-2 of BASELINE.json, and with each other far tighter."""
    import quant
    R, C, group = 512, 1024, 128
    g = torch.Generator().manual_seed(M)
    G = C // group
    meta = {'qweight': torch.randint(0, 16, (R, C), generator=g, dtype=torch.uint8),
            'scale': (torch.rand(R, G, generator=g) * 0.01 + 1e-3).half(),
            'qzero': torch.randint(0, 16, (R, G), generator=g, dtype=torch.uint8),
            'perm_inv': torch.randperm(C, generator=g).to(torch.int16),
            'group_sizes': torch.full([G], group, dtype=torch.int16), 'qscale': None, 'sscale': None,
            'group_bit_widths': torch.full([G], 4, dtype=torch.uint8), 'scale_bit_width': None}
    lin = quant.W4A16Linear.from_quant_meta(meta, torch.device('cuda'))
    assert M >= lin.dense_rows(R, C)
    x = torch.randn(M, C, generator=g).half().cuda()
    y_dense_route = lin(x).float()
    lin.dense_route = False
    y_fused = lin(x).float()
    ref = quant.reconstruct_nn_linear(meta, torch.float16, torch.device('cuda'))
    want = ref(x).float()
    for y in (y_dense_route, y_fused):
        assert ((y - want).norm() / want.norm()).item() <= 1e-2
    assert ((y_dense_route - want).norm() / want.norm()).item() <= 2e-3     # same 16-bit weights, fp32 accumulation


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
@pytest.mark.parametrize('M,N,K', [(1, 256, 64), (128, 256, 128), (300, 520, 200), (257, 1000, 4096), (2048, 4096, 4096),
                                   (513, 11008, 4096), (96, 4096, 11008)])
def test_dense_gemm_nt_vs_fp32_reference(M, N, K, dtype):
    """gq_gemm_f16_nt (the GEMM of K6's large-batch route): y = x w^T against a float32 matmul of the same 16-bit operands;
    ragged M / N / K edges (TMA zero fill, partial stores) included.  Tolerance: one 16-bit rounding of y plus fp32
    accumulation-order noise."""
    import gptq
    g = torch.Generator(device='cuda').manual_seed(M + N + K)
    x = torch.randn(M, K, device='cuda', generator=g).to(dtype)
    w = (torch.randn(N, K, device='cuda', generator=g) * 0.05).to(dtype)
    y = gptq.gemm_f16_nt(x, w)
    want = x.float() @ w.float().t()
    assert y.shape == (M, N) and y.dtype == dtype
    eps = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    err = (y.float() - want).abs()
    bound = eps * want.abs() + 1e-4 * (x.float().abs() @ w.float().abs().t())
    assert bool((err <= bound).all()), float((err / bound.clamp_min(1e-30)).max())


def test_w4a16_linear_large_batch_route_uses_own_gemm():
    """W4A16Linear above its dense_rows threshold: one dequantisation + gq_gemm_f16_nt, no torch.matmul; result within 1e-2 of
    the reference's dense forward (a8)."""
    import gptq
    from quant import W4A16Linear, pack_quant_meta, reconstruct_nn_linear
    R, C = 512, 1024
    g = torch.Generator().manual_seed(4)
    meta = {'qweight': torch.randint(0, 16, (R, C), generator=g, dtype=torch.uint8),
            'scale': (torch.rand(R, C // 128, generator=g) * 0.01 + 1e-3).half(),
            'qzero': torch.randint(0, 16, (R, C // 128), generator=g, dtype=torch.uint8),
            'qscale': None, 'sscale': None, 'perm_inv': torch.randperm(C, generator=g).to(torch.int16),
            'group_sizes': torch.full((C // 128,), 128, dtype=torch.int16),
            'group_bit_widths': torch.full((C // 128,), 4, dtype=torch.uint8), 'scale_bit_width': None}
    lin = W4A16Linear(pack_quant_meta(meta, torch.device('cuda')))
    ref = reconstruct_nn_linear(meta, torch.float16, torch.device('cuda'))
    x = torch.randn(2, 700, C, device='cuda', dtype=torch.float16)
    n0 = gptq.lib.gq_launch_count()
    y = lin(x)
    assert gptq.lib.gq_launch_count() - n0 >= 2                  # dequant + GEMM are this library's launches
    want = ref(x)
    assert float((y.float() - want.float()).norm() / want.float().norm()) <= 1e-2
