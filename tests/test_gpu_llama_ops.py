"""Row f1 (driver costs): the one-pass elementwise kernels of csrc/llama_ops.cu against the PyTorch expressions of the
decoder block they replace in llama_driver.py (the HF Llama forms the reference's quantize_llama replays, llama.py:121-125,
176-189).  Rotary embedding and silu * up are evaluated operation by operation with the same roundings, so they must be
bit-equal; rms_norm's fp32 mean is summed in a different order, so a result may differ in its last place (by at most two 16-bit ulps: two roundings) in rare elements."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DTYPES = [torch.float16, torch.bfloat16]


def _ulp_diff(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """distance in units of the 16-bit format's last place (same-sign finite values)"""
    ia = a.contiguous().view(torch.int16).to(torch.int32)
    ib = b.contiguous().view(torch.int16).to(torch.int32)
    return (ia - ib).abs()


@pytest.mark.parametrize('dtype', DTYPES)
@pytest.mark.parametrize('rows,D', [(7, 64), (1000, 4096), (33, 5120), (16, 8192), (5, 1000)])
def test_rms_norm_matches_the_pytorch_expression(rows, D, dtype):
    import gptq
    import llama_driver as LD
    g = torch.Generator().manual_seed(rows + D)
    x = (torch.randn(rows, D, generator=g) * 3 + 0.5).to(dtype).cuda()
    w = (torch.rand(D, generator=g) + 0.5).to(dtype).cuda()
    xf = x.float()
    want = w * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)).to(dtype)
    got = gptq.llama_rms_norm(x, w, 1e-5)
    out = torch.empty_like(x)
    assert LD.rms_norm(x.view(1, rows, D), w, 1e-5, out=out.view(1, rows, D)).data_ptr() == out.data_ptr()
    inplace = gptq.llama_rms_norm(x.clone(), w, 1e-5)
    torch.cuda.synchronize()
    d = _ulp_diff(got, want)       # a flipped first rounding can move the second one too: at most 2 ulps, and rarely
    assert int(d.max()) <= 2 and float((d == 0).float().mean()) >= 0.999
    assert torch.equal(out, got) and torch.equal(inplace, got)
    # against float64 on the same inputs: within the two roundings of the expression
    ref = (w.double() * (x.double() * torch.rsqrt(x.double().pow(2).mean(-1, keepdim=True) + 1e-5)))
    eps = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    assert ((got.double() - ref).abs() <= 1.01 * eps * ref.abs() + 1e-6).all()


@pytest.mark.parametrize('dtype', DTYPES)
@pytest.mark.parametrize('B,N,heads,hd', [(2, 64, 4, 128), (1, 2048, 32, 128), (3, 17, 2, 64), (2, 8, 1, 16)])
def test_rope_in_place_is_bit_equal_to_rotate_half(B, N, heads, hd, dtype):
    import gptq
    import llama_driver as LD
    g = torch.Generator().manual_seed(B + N + heads + hd)
    x = torch.randn(B, N, heads * hd, generator=g).to(dtype).cuda()
    cos, sin = LD.rope_tables(N, hd, 10000.0, x.device, dtype)
    want = LD.apply_rope(x.view(B, N, heads, hd).transpose(1, 2), cos, sin).transpose(1, 2).reshape(B, N, heads * hd)
    got = gptq.llama_rope_(x.clone(), cos, sin, heads)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


@pytest.mark.parametrize('dtype', DTYPES)
@pytest.mark.parametrize('shape', [(3, 8), (16, 2048, 88), (4, 11008)])
def test_silu_mul_matches_f_silu_times_up(shape, dtype):
    import gptq
    g = torch.Generator().manual_seed(sum(shape))
    a = (torch.randn(*shape, generator=g) * 4).to(dtype).cuda()
    b = torch.randn(*shape, generator=g).to(dtype).cuda()
    want = F.silu(a) * b
    got = gptq.llama_silu_mul(a, b)
    out = torch.empty_like(a)
    gptq.llama_silu_mul(a, b, out=out)
    torch.cuda.synchronize()
    d = _ulp_diff(got, want)
    assert int(d.max()) <= 1 and float((d == 0).float().mean()) >= 0.9999
    assert torch.equal(out, got)


def test_attention_with_fused_rope_equals_the_pytorch_route():
    """llama_driver.attention on CUDA (rotary in place, result into `out`) == the same function's PyTorch route (fp32 inputs
    are not taken by the kernels) evaluated on the 16-bit values."""
    import llama_driver as LD
    B, N, heads, hd = 2, 128, 4, 64
    g = torch.Generator().manual_seed(5)
    q, k, v = (torch.randn(B, N, heads * hd, generator=g).half().cuda() for _ in range(3))
    cos, sin = LD.rope_tables(N, hd, 10000.0, q.device, torch.float16)
    qr = LD.apply_rope(q.view(B, N, heads, hd).transpose(1, 2), cos, sin)
    kr = LD.apply_rope(k.view(B, N, heads, hd).transpose(1, 2), cos, sin)
    want = F.scaled_dot_product_attention(qr, kr, v.view(B, N, heads, hd).transpose(1, 2), is_causal=True).transpose(1, 2).reshape(B, N, heads * hd)
    out = torch.empty_like(q)
    LD.attention(q.clone(), k.clone(), v, heads, cos, sin, out=out)
    torch.cuda.synchronize()
    assert torch.allclose(out.float(), want.float(), atol=2e-3, rtol=2e-3)


def test_llama_ops_reject_what_they_do_not_take():
    import gptq
    x = torch.randn(4, 16384, device='cuda').half()
    with pytest.raises(gptq.GptqB200Error):
        gptq.llama_rms_norm(x, torch.ones(16384, device='cuda').half(), 1e-5)       # rows longer than 8192
    with pytest.raises(gptq.GptqB200Error):
        gptq.llama_silu_mul(x, x.float())
    # empty inputs are no-ops
    e = torch.empty(0, 64, device='cuda').half()
    assert gptq.llama_rms_norm(e, torch.ones(64, device='cuda').half(), 1e-5).shape == (0, 64)
    assert gptq.llama_silu_mul(e, e).shape == (0, 64)
