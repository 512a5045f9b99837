"""Layer-by-layer driver (gptq-v2_b200/llama_driver.py) against transformers' own Llama and the reference flow's
invariants (llama.py:65-204): block forward parity, result layout, and -- on the GPU -- that the hidden state the
driver carries is the hidden state of the QUANTISED model (replay through reconstructed weights)."""
import pytest
import torch

import llama_driver as L


def _tiny_hf(layers=2, hidden=128, inter=256, heads=4, vocab=500, seed=0):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(hidden_size=hidden, intermediate_size=inter, num_hidden_layers=layers, num_attention_heads=heads,
                      num_key_value_heads=heads, vocab_size=vocab, max_position_embeddings=256)
    m = LlamaForCausalLM(cfg).eval()
    for p in m.parameters():
        if p.dim() == 1:
            p.data = 1 + 0.1 * torch.randn_like(p)
    return m, cfg


def test_block_forward_matches_transformers_cpu():
    m, cfg = _tiny_hf()
    m = m.float()
    tok = torch.randint(0, cfg.vocab_size, (3, 64))
    with torch.no_grad():
        hs = m.model(tok, output_hidden_states=True).hidden_states
    shape = L.LlamaShape(cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers, cfg.num_attention_heads,
                         cfg.vocab_size, cfg.rms_norm_eps, L._rope_theta(cfg))
    x = hs[0]
    for i in range(cfg.num_hidden_layers):
        x = L.block_forward(shape, L.hf_block_weights(m.model.layers[i], torch.device('cpu')), x)
        if i + 1 < cfg.num_hidden_layers:                       # the last entry has the final norm applied
            assert torch.allclose(x, hs[i + 1], atol=1e-5, rtol=1e-5)
    assert torch.allclose(L.rms_norm(x, m.model.norm.weight, cfg.rms_norm_eps), hs[-1], atol=1e-5, rtol=1e-5)


def test_random_block_weights_deterministic_cpu():
    shape = L.LlamaShape(hidden=64, inter=128, layers=2, heads=2, vocab=100)
    a = L.random_block_weights(shape, 1, torch.device('cpu'), seed=3)
    b = L.random_block_weights(shape, 1, torch.device('cpu'), seed=3)
    c = L.random_block_weights(shape, 0, torch.device('cpu'), seed=3)
    assert set(a) == set(L.LINEAR_NAMES) | {'input_layernorm', 'post_attention_layernorm'}
    assert all(torch.equal(a[k], b[k]) for k in a)
    assert not torch.equal(a['mlp.down_proj'], c['mlp.down_proj'])
    assert a['mlp.down_proj'].shape == (64, 128) and a['mlp.up_proj'].shape == (128, 64)


@pytest.mark.gpu
@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
def test_quantize_hf_llama_replays_quantised_model(dtype):
    from quant import reconstruct_nn_linear
    dev = torch.device('cuda:0')
    m, cfg = _tiny_hf(layers=2, hidden=256, inter=512, heads=4)
    m = m.to(dtype).to(dev)
    g = torch.Generator().manual_seed(5)
    tok = torch.randint(0, cfg.vocab_size, (24, 128), generator=g)
    res = L.quantize_hf_llama(m, tok, dev, batch_size=16)       # ragged last batch (24 = 16 + 8)
    names = [f'model.layers.{i}.{n}' for i in range(2) for n in L.LINEAR_NAMES]
    assert list(res['data']) == names and list(res['metrics']) == names
    for n in names:
        meta = res['data'][n]
        assert set(meta) == {'qweight', 'scale', 'qzero', 'qscale', 'sscale', 'perm_inv', 'group_sizes',
                             'group_bit_widths', 'scale_bit_width'}                      # gptq_py.py:239-249
        assert meta['qweight'].device.type == 'cpu' and meta['qweight'].dtype == torch.uint8
        assert int(meta['qweight'].max()) <= 15
        met = res['metrics'][n]
        assert set(met) == {'gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'}
        assert all(v == v and v >= 0 for v in met.values())
        assert met['l2.0_error'] < 0.05 * met['l2.0_norm']                              # 4-bit g128 sanity
    # swap the reconstructed linears into the HF model: its hidden states must be what the driver carried forward
    for n in names:
        lin = reconstruct_nn_linear(res['data'][n], dtype=dtype, device=dev)
        m.get_submodule(n).weight.data = lin.weight.data
    with torch.no_grad():
        hs = m.model(tok.to(dev), output_hidden_states=True).hidden_states
    x = L.block_forward(L.LlamaShape(256, 512, 2, 4, cfg.vocab_size, cfg.rms_norm_eps, L._rope_theta(cfg)),
                        L.hf_block_weights(m.model.layers[0], dev), hs[0])
    final = res['final_hidden']
    ref_final = L.block_forward(L.LlamaShape(256, 512, 2, 4, cfg.vocab_size, cfg.rms_norm_eps, L._rope_theta(cfg)),
                                L.hf_block_weights(m.model.layers[1], dev), x)
    tol = 2e-2 if dtype == torch.float16 else 1e-1
    assert torch.allclose(final.float(), ref_final.float(), atol=tol, rtol=tol)
    assert torch.allclose(x.float(), hs[1].float(), atol=tol, rtol=tol)
    assert set(res['timing']) >= {'hessian', 'invert', 'solve', 'forward'}


@pytest.mark.gpu
def test_quantize_llama_random_init_gptq_beats_rtn():
    """GPTQ's purpose on this path: the layer OUTPUT error on the calibration inputs is lower than round-to-nearest's
    at the same grid (the solver compensates along H^-1)."""
    from gptq_py import gptq_quant
    from quant import reconstruct_nn_linear
    dev = torch.device('cuda:0')
    shape = L.LlamaShape(hidden=256, inter=512, layers=1, heads=4, vocab=300)
    inps = L.synthetic_inputs(shape, 32, 128, dev)
    # correlated features make the Hessian non-trivial
    mix = torch.randn(256, 256, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).half() * 0.1
    inps = inps + inps @ mix
    w = L.random_block_weights(shape, 0, dev)
    res = L.quantize_llama(shape, lambda i: w, inps, batch_size=16)
    h1 = L.rms_norm(inps, w['input_layernorm'], shape.rms_eps).float()
    W = w['self_attn.q_proj']
    G = W.shape[1] // 128
    gs = torch.full([G], 128, dtype=torch.int32, device=dev)
    gb = torch.full([G], 4, dtype=torch.int32, device=dev)
    rtn = gptq_quant(W, None, gs, gb, quant_mse=False, save_device=dev)
    w_rtn = reconstruct_nn_linear(rtn['quant_meta'], torch.float16, dev).weight.float()
    w_gptq = reconstruct_nn_linear(res['data']['model.layers.0.self_attn.q_proj'], torch.float16, dev).weight.float()
    ref = h1 @ W.float().t()
    e_rtn = (h1 @ w_rtn.t() - ref).pow(2).mean().item()
    e_gptq = (h1 @ w_gptq.t() - ref).pow(2).mean().item()
    assert e_gptq < e_rtn


@pytest.mark.gpu
@pytest.mark.parametrize('opts', [{'quant_mse': True}, {'quant_symmetric': True}, {'quant_mse': True, 'scale_bit_width': 4}])
def test_quantize_llama_with_quantizer_options(opts):
    """The driver hands the SURVEY 8f rank-4 Quantizer options through to gptq_quant: the block still replays through its
    quantised weights, EXL2 metadata reconstructs, and the MSE search does not make the weight error worse than the plain
    grid on round-to-nearest."""
    from gptq_py import gptq_quant
    from quant import reconstruct_nn_linear
    dev = torch.device('cuda:0')
    shape = L.LlamaShape(hidden=256, inter=512, layers=1, heads=4, vocab=300)
    inps = L.synthetic_inputs(shape, 16, 128, dev)
    w = L.random_block_weights(shape, 0, dev)
    o = dict(opts)
    if 'scale_bit_width' in o:
        o['scale_bit_width'] = torch.tensor(o['scale_bit_width'])
    res = L.quantize_llama(shape, lambda i: w, inps, batch_size=8, quant_options=o)
    plain = L.quantize_llama(shape, lambda i: w, inps, batch_size=8)
    assert set(res['data']) == set(plain['data'])
    for name, meta in res['data'].items():
        what = reconstruct_nn_linear(meta, torch.float16, dev).weight
        assert torch.isfinite(what).all()
        if 'scale_bit_width' in opts:
            assert meta['qscale'] is not None and meta['sscale'].shape == (meta['scale'].shape[1],)
            assert int(meta['qscale'].min()) >= 1 and int(meta['qscale'].max()) <= 16
        rel = res['metrics'][name]['l2.0_error'] / res['metrics'][name]['l2.0_norm']
        assert rel < 0.05, (name, rel)                       # 4-bit g128: a few percent of the weight energy at most
    assert torch.isfinite(res['final_hidden']).all()
    if opts.get('quant_mse') and 'scale_bit_width' not in opts:
        W = w['mlp.down_proj']
        G = W.shape[1] // 128
        gs = torch.full([G], 128, dtype=torch.int32, device=dev)
        gb = torch.full([G], 4, dtype=torch.int32, device=dev)
        e0 = gptq_quant(W, None, gs, gb, quant_mse=False, save_device=dev)['metrics']['l2.0_error']
        e1 = gptq_quant(W, None, gs, gb, quant_mse=True, quant_norm=2.0, save_device=dev)['metrics']['l2.0_error']
        assert e1 <= e0 * (1 + 1e-6)                         # the grid contains p = 1, and norm 2 is the metric itself
