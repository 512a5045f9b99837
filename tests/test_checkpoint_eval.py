"""packed on-disk model (checkpoint.py) and layer-streamed perplexity on the
quantised model through K6 (llama_driver.evaluate_llama, the reference's llama.py:207-247)."""
import copy

import pytest
import torch

import llama_driver as L


def _tiny_hf(layers=2, hidden=256, inter=512, heads=4, vocab=500, seed=0):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(hidden_size=hidden, intermediate_size=inter, num_hidden_layers=layers, num_attention_heads=heads,
                      num_key_value_heads=heads, vocab_size=vocab, max_position_embeddings=256)
    m = LlamaForCausalLM(cfg).eval()
    for p in m.parameters():
        if p.dim() == 1:
            p.data = 1 + 0.1 * torch.randn_like(p)
    return m, cfg


def _shape(cfg):
    return L.LlamaShape(cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers, cfg.num_attention_heads,
                        cfg.vocab_size, cfg.rms_norm_eps, L._rope_theta(cfg))


def _reference_ppl(model, tok, batch_size):
    """llama.py:233-242 on a transformers model: mean over sequences of the per-batch mean token NLL, exponentiated."""
    nlls = []
    with torch.no_grad():
        for i in range(0, len(tok), batch_size):
            logits = model(tok[i:i + batch_size]).logits
            nll = torch.nn.functional.cross_entropy(logits[:, :-1].flatten(end_dim=-2).float(), tok[i:i + batch_size, 1:].flatten())
            nlls.extend([nll] * logits.shape[0])
    return torch.stack(nlls).float().mean().exp()


def test_evaluate_llama_matches_transformers_cpu():
    m, cfg = _tiny_hf(hidden=128, inter=256)
    m = m.float()
    tok = torch.randint(0, cfg.vocab_size, (5, 48), generator=torch.Generator().manual_seed(3))
    dev = torch.device('cpu')
    ppl = L.evaluate_llama(_shape(cfg), lambda i: L.hf_block_weights(m.model.layers[i], dev), m.model.embed_tokens.weight.detach(),
                           m.model.norm.weight.detach(), m.lm_head.weight.detach(), tok, batch_size=2)
    want = _reference_ppl(m, tok, 2)
    assert abs(float(ppl) - float(want)) <= 1e-4 * float(want)


def test_load_packed_rejects_other_files_cpu(tmp_path):
    import checkpoint
    p = tmp_path / 'x.pt'
    torch.save({'data': {}, 'metrics': {}}, p)                       # the reference's own result dict is not a packed file
    with pytest.raises(ValueError):
        checkpoint.load_packed(str(p))
    with pytest.raises(ValueError):
        checkpoint.save_packed({'layers': {}}, str(tmp_path / 'y.pt'))


@pytest.mark.gpu
def test_packed_roundtrip_and_quantised_model_eval(tmp_path):
    import checkpoint
    from quant import reconstruct_nn_linear
    dev = torch.device('cuda:0')
    dtype = torch.float16
    m, cfg = _tiny_hf()
    m = m.to(dtype).to(dev)
    shape = _shape(cfg)
    g = torch.Generator().manual_seed(11)
    calib = torch.randint(0, cfg.vocab_size, (16, 128), generator=g)
    res = L.quantize_hf_llama(m, calib, dev, batch_size=8)

    # ---- rank 2: pack -> save -> load -> unpack is the reference's quant_meta, bit for bit, at about half the bytes
    packed = checkpoint.pack_results(res, dev)
    path = tmp_path / 'model.gptq_b200.pt'
    checkpoint.save_packed(packed, str(path))
    loaded = checkpoint.load_packed(str(path))
    assert list(loaded['layers']) == list(res['data']) and loaded['metrics'] == res['metrics']
    raw = 0
    for name, meta in res['data'].items():
        back = checkpoint.unpack_layer(loaded['layers'][name], dev)
        assert set(back) == set(meta)
        for k, v in meta.items():
            if v is None:
                assert back[k] is None
            else:
                assert back[k].dtype == v.dtype and torch.equal(back[k], v), (name, k)
        raw += sum(v.numel() * v.element_size() for v in meta.values() if torch.is_tensor(v))
    assert checkpoint.packed_bytes(loaded) < 0.62 * raw

    # ---- the packed model is runnable: W4A16Linear modules inside the transformers model
    tok = torch.randint(0, cfg.vocab_size, (6, 96), generator=g).to(dev)
    dense = copy.deepcopy(m)
    for name, meta in res['data'].items():
        dense.get_submodule(name).weight.data = reconstruct_nn_linear(meta, dtype=dtype, device=dev).weight.data
    q = checkpoint.apply_to_model(copy.deepcopy(m), loaded, dev, dtype)
    with torch.no_grad():
        want = dense(tok).logits.float()
        got = q(tok).logits.float()
    assert ((got - want).norm() / want.norm()).item() <= 1e-2

    # ---- rank 3: perplexity of the quantised model straight from packed codes (K6) vs its dense reconstruction
    norms = lambda i: {'input_layernorm': m.model.layers[i].input_layernorm.weight.detach(),           # noqa: E731
                       'post_attention_layernorm': m.model.layers[i].post_attention_layernorm.weight.detach()}
    args = (m.model.embed_tokens.weight.detach(), m.model.norm.weight.detach(), m.lm_head.weight.detach(), tok.cpu())
    ppl_packed = float(L.evaluate_llama(shape, L.packed_block_weights(loaded, norms, dev, dtype), *args, batch_size=2))
    ppl_dense = float(L.evaluate_llama(shape, lambda i: L.hf_block_weights(dense.model.layers[i], dev), *args, batch_size=2))
    ppl_hf = float(_reference_ppl(dense, tok, 2))
    assert abs(ppl_dense - ppl_hf) <= 2e-2 * ppl_hf
    assert abs(ppl_packed - ppl_dense) <= 1e-2 * ppl_dense


@pytest.mark.gpu
@pytest.mark.parametrize('bits', [3, (4, 3, 2, 4)])
def test_packed_roundtrip_3bit_and_mixed_widths(bits, tmp_path):
    """BASELINE configs[3] is 3-bit: its quant_meta packs (one code per nibble), survives save / load bit for bit with
    its group_bit_widths, and K6 evaluates it like the dense reconstruction (quant.py:262-279)."""
    import checkpoint
    from gptq_py import HessianHook, gptq_quant
    from quant import reconstruct_nn_linear
    dev = torch.device('cuda:0')
    g = torch.Generator().manual_seed(5)
    R, C, gs = 256, 512, 128
    W = (torch.randn(R, C, generator=g) * 0.02).half().to(dev)
    x = torch.randn(4, 1024, C, generator=g).half().to(dev)
    hook = HessianHook()
    hook.add_batch(x)
    hook.invert()
    G = C // gs
    widths = torch.full([G], bits, dtype=torch.uint8) if isinstance(bits, int) else torch.tensor(bits, dtype=torch.uint8)
    res = gptq_quant(W, hook, torch.full([G], gs, dtype=torch.int16), widths, gptq_use_kernel=False,
                     gptq_block_sizes=torch.full([G], gs, dtype=torch.int16), quant_mse=False)
    meta = res['quant_meta']
    assert int(meta['qweight'].max()) <= 2 ** int(widths.max()) - 1
    packed = checkpoint.pack_results({'data': {'lin': meta}, 'metrics': {'lin': res['metrics']}}, dev)
    path = tmp_path / 'm.pt'
    checkpoint.save_packed(packed, str(path))
    loaded = checkpoint.load_packed(str(path))
    back = checkpoint.unpack_layer(loaded['layers']['lin'], dev)
    for k, v in meta.items():
        if v is None:
            assert back[k] is None
        else:
            assert back[k].dtype == v.dtype and torch.equal(back[k].cpu(), v.cpu()), k
    lin = checkpoint.quantized_linears(loaded, dev, torch.float16)['lin']
    dense = reconstruct_nn_linear(meta, dtype=torch.float16, device=dev)
    xe = torch.randn(5, C, generator=g).half().to(dev)
    with torch.no_grad():
        want, got = dense(xe).float(), lin(xe).float()
    assert ((got - want).norm() / want.norm()).item() <= 1e-2
