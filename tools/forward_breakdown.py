#!/usr/bin/env python
"""Where the replay forward of one Llama-2-7B decoder block spends its time (llama_driver's functions, 128 x 2048 tokens in
batches of 16 sequences, dense fp16 weights): CUDA events around each component, summed over the batches, one JSON line."""
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch
import torch.nn.functional as F
import llama_driver as LD

dev = torch.device('cuda')
shape = LD.LlamaShape.llama2_7b()
n_seq, N, bs = 128, 2048, 16
w = LD.random_block_weights(shape, 0, dev)
x_all = (torch.randn(n_seq, N, shape.hidden, device=dev) * 0.5).half()
cos, sin = LD.rope_tables(N, shape.hidden // shape.heads, shape.rope_theta, dev, torch.float16)
spans = {}


class span:
    def __init__(self, name): self.name = name
    def __enter__(self):
        self.a = torch.cuda.Event(enable_timing=True); self.b = torch.cuda.Event(enable_timing=True); self.a.record()
    def __exit__(self, *e):
        self.b.record(); spans.setdefault(self.name, []).append((self.a, self.b))


def one_pass():
    h1 = torch.empty_like(x_all); att = torch.empty_like(x_all); h2 = torch.empty_like(x_all); h3 = torch.empty_like(x_all)
    act = torch.empty(n_seq, N, shape.inter, dtype=torch.float16, device=dev)
    for i in range(0, n_seq, bs):
        b = slice(i, i + bs)
        with span('rms_norm x2'):
            LD.rms_norm(x_all[b], w['input_layernorm'], shape.rms_eps, out=h1[b])
        with span('q k v GEMMs'):
            q, k, v = LD.linear(h1[b], w['self_attn.q_proj']), LD.linear(h1[b], w['self_attn.k_proj']), LD.linear(h1[b], w['self_attn.v_proj'])
        with span('rope + attention + transposing copy'):
            LD.attention(q, k, v, shape.heads, cos, sin, out=att[b])
        with span('o GEMM'):
            o = LD.linear(att[b], w['self_attn.o_proj'])
        with span('residual adds'):
            h2[b] = x_all[b] + o
        with span('rms_norm x2'):
            LD.rms_norm(h2[b], w['post_attention_layernorm'], shape.rms_eps, out=h3[b])
        with span('gate up GEMMs'):
            g, u = LD.linear(h3[b], w['mlp.gate_proj']), LD.linear(h3[b], w['mlp.up_proj'])
        with span('silu * up'):
            LD.silu_mul(g, u, out=act[b])
        with span('down GEMM'):
            d = LD.linear(act[b], w['mlp.down_proj'])
        with span('residual adds'):
            h2[b] += d
    return h2


for it in range(3):
    spans.clear()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); one_pass(); e1.record()
    torch.cuda.synchronize()
tok = n_seq * N
flops = {'q k v GEMMs': 3 * 2 * tok * shape.hidden ** 2, 'o GEMM': 2 * tok * shape.hidden ** 2,
         'gate up GEMMs': 2 * 2 * tok * shape.hidden * shape.inter, 'down GEMM': 2 * tok * shape.hidden * shape.inter}
out = {'what': 'replay forward of one 7B block, ms per component (last of 3 passes)', 'total_ms': e0.elapsed_time(e1)}
for name, evs in spans.items():
    ms = sum(a.elapsed_time(b) for a, b in evs)
    out[name] = round(ms, 2) if name not in flops else {'ms': round(ms, 2), 'tflops': round(flops[name] / ms / 1e9, 1)}
print(json.dumps(out))
