"""gq_w4a16_gemv_static: the allocation-free batch 1-2 variant of gq_w4a16_gemm (include/gptq_b200.h).

No counterpart in the reference; added for callers that cannot call cudaMalloc.  Called through the C ABI (ctypes) and
checked against the existing entry point gq_w4a16_gemm on the same operands:
  * against its CUDA-core route (gq_set_w4a16_path(1): the same kernels with a caller-provided workspace): bit for bit;
  * against its default route (the warp-level mma kernel): rel <= 5e-3, the bound test_w4a16_paths_agree uses;
on several calls in a row that share the static scratch (different M, K, dtype, permutation), with the device's free
memory unchanged across a call.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

GQ_ERR_LIMIT = 9


def _problem(M, N, K, group, dtype, use_perm, seed):
    g = torch.Generator().manual_seed(seed)
    G = K // group
    codes = torch.randint(0, 16, (N, K), generator=g, dtype=torch.uint8)
    scale_t = (torch.rand(G, N, generator=g) * 0.01 + 1e-3).to(dtype).cuda()
    qzero_t = torch.randint(0, 16, (G, N), generator=g, dtype=torch.uint8).cuda()
    perm = torch.randperm(K, generator=g).to(torch.int32).cuda() if use_perm else None
    x = torch.randn(M, K, generator=g).to(dtype).cuda()
    return codes, scale_t, qzero_t, perm, x


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _static_call(lib, x, packed, scale_t, qzero_t, perm, y, group):
    from gptq import _cabi
    M, K = x.shape
    N = packed.shape[0]
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.gq_w4a16_gemv_static(_p(x), _p(packed), _p(scale_t), _p(qzero_t), _p(perm), _p(y), M, N, K, group,
                                    _cabi.DTYPE_CODE[x.dtype], stream)


def _existing_call(lib, x, packed, scale_t, qzero_t, perm, y, group, path):
    """gq_w4a16_gemm with a caller-provided workspace, on the route `path` (0 default, 1 CUDA-core GEMV for M <= 2)."""
    from gptq import _cabi
    M, K = x.shape
    N = packed.shape[0]
    work = torch.empty(max(int(lib.gq_w4a16_workspace_bytes(M, N, K)), 256), dtype=torch.uint8, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    old = lib.gq_set_w4a16_path(path)
    try:
        return lib.gq_w4a16_gemm(_p(x), _p(packed), _p(scale_t), _p(qzero_t), _p(perm), _p(y), _p(work), M, N, K, group,
                                 _cabi.DTYPE_CODE[x.dtype], stream)
    finally:
        lib.gq_set_w4a16_path(old)


CALLS = [
    # M, N, K, group, dtype, use_perm  -- consecutive calls share the scratch: larger, then smaller, then larger again
    (2, 4096, 4096, 128, torch.float16, True),
    (1, 1024, 11008, 128, torch.float16, True),
    (1, 592, 384, 128, torch.bfloat16, False),
    (2, 256, 16384, 64, torch.bfloat16, True),
    (2, 4096, 4096, 128, torch.float16, True),
]


def test_static_scratch_variant_matches_existing_entry_point():
    import gptq
    from gptq import _cabi
    lib = _cabi.lib
    for i, (M, N, K, group, dtype, use_perm) in enumerate(CALLS):
        codes, scale_t, qzero_t, perm, x = _problem(M, N, K, group, dtype, use_perm, seed=1000 + i)
        packed = gptq.pack_int4(codes.cuda())
        y_new = torch.full((M, N), float('nan'), dtype=dtype, device='cuda')
        y_gemv = torch.empty_like(y_new)
        y_def = torch.empty_like(y_new)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        assert _static_call(lib, x, packed, scale_t, qzero_t, perm, y_new, group) == 0
        torch.cuda.synchronize()
        if i > 0:       # the first call of a process may page kernels in; later ones must leave device memory alone
            assert torch.cuda.mem_get_info()[0] >= free_before
        assert _existing_call(lib, x, packed, scale_t, qzero_t, perm, y_gemv, group, path=1) == 0
        assert _existing_call(lib, x, packed, scale_t, qzero_t, perm, y_def, group, path=0) == 0
        torch.cuda.synchronize()
        assert torch.isfinite(y_new.float()).all()
        assert torch.equal(y_new, y_gemv), (i, (y_new.float() - y_gemv.float()).abs().max().item())
        rel = ((y_new.float() - y_def.float()).norm() / y_def.float().norm()).item()
        assert rel <= 5e-3, (i, rel)


def test_static_scratch_variant_through_python_op_twice():
    """The package-level op (caller-provided `out`), two calls in a row on the same operands and on new ones."""
    import gptq
    outs = []
    for seed in (7, 7, 8):
        codes, scale_t, qzero_t, perm, x = _problem(1, 2048, 4096, 128, torch.float16, True, seed)
        packed = gptq.pack_int4(codes.cuda())
        out = torch.empty(1, 2048, dtype=torch.float16, device='cuda')
        gptq.w4a16_gemv_static(x, packed, scale_t, qzero_t, perm, out, 128)
        want = gptq.w4a16_gemm(x, packed, scale_t, qzero_t, perm, 128)
        torch.cuda.synchronize()
        assert ((out.float() - want.float()).norm() / want.float().norm()).item() <= 5e-3
        outs.append(out)
    assert torch.equal(outs[0], outs[1])
    assert not torch.equal(outs[0], outs[2])


def test_static_scratch_variant_rejects_sizes_beyond_the_scratch():
    import gptq
    from gptq import _cabi
    lib = _cabi.lib
    codes, scale_t, qzero_t, perm, x = _problem(3, 64, 256, 128, torch.float16, False, 1)
    packed = gptq.pack_int4(codes.cuda())
    y = torch.zeros(3, 64, dtype=torch.float16, device='cuda')
    assert _static_call(lib, x, packed, scale_t, qzero_t, None, y, 128) == GQ_ERR_LIMIT          # M > 2
    K = 16384 + 128
    codes, scale_t, qzero_t, perm, x = _problem(1, 32, K, 128, torch.float16, False, 2)
    packed = gptq.pack_int4(codes.cuda())
    y1 = torch.zeros(1, 32, dtype=torch.float16, device='cuda')
    assert _static_call(lib, x, packed, scale_t, qzero_t, None, y1, 128) == GQ_ERR_LIMIT         # K > 16384
    torch.cuda.synchronize()
    assert float(y.abs().sum()) == 0.0 and float(y1.abs().sum()) == 0.0                          # nothing was launched
    assert b'allocation-free' in lib.gq_error_string(GQ_ERR_LIMIT)
    with pytest.raises(gptq.GptqB200Error):
        gptq.w4a16_gemv_static(x, packed, scale_t, qzero_t, None, y1, 128)
    # an empty batch is a no-op, as in gq_w4a16_gemm
    assert lib.gq_w4a16_gemv_static(None, None, None, None, None, None, 0, 32, 256, 128, 0, None) == 0
