"""On-disk form of a quantised model and the way back into a runnable one  (SURVEY.md section 8f rank 2).

The reference saves `torch.save(results, path)` with results = {'data': {name: quant_meta}, 'metrics': {...}}
(/root/reference/main.py:37-38, llama.py:82-85,173): one uint8 per 4-bit code.  This module keeps that dict readable
and adds the packed form K6 consumes:

    pack_results(results)           reference-format results -> {'format', 'layers': {name: packed layer}, 'metrics'}
    save_packed / load_packed       torch.save / torch.load of that dict (tensors on the CPU)
    unpack_layer(layer)             packed layer -> the reference's 9-key quant_meta, bit for bit
    quantized_linears(packed, dev)  {name: W4A16Linear}
    apply_to_model(model, packed)   swap the nn.Linear modules of a transformers Llama for W4A16Linear in place

A packed layer holds: qweight_packed (R, C/8) int32 (column 8k+i in bits [4i, 4i+4) of word k, act-order column order),
scale_t (G, R) 16-bit, qzero_t (G, R) uint8 (group-major, K6's layout), perm (C) int32 (quant position -> input
column) or None, group_size, bits (None + group_bit_widths (G) uint8 when the widths are mixed), shape (R, C): 0.5 B per
weight + 3 B per group instead of 1 B per weight.  Codes of 2 / 3 bits are kept one per nibble like 4-bit ones.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

import gptq as _ops
from quant import W4A16Linear, pack_quant_meta

FORMAT = 'gptq_b200.packed.v1'


def pack_layer(quant_meta: dict, device: torch.device) -> dict:
    """One quant_meta (gptq_py.py:239-249) -> packed layer with CPU tensors.  Codes of up to 4 bits, one per nibble: 4-bit
    (llama.py:147), 3-bit (BASELINE configs[3]) and mixed per-group widths (their widths ride along in `group_bit_widths`;
    K6 does not need them: a code is just a number below 16)."""
    bits = quant_meta['group_bit_widths']
    if int(bits.max()) > 4:
        raise NotImplementedError('packed storage holds one code per nibble: bit widths up to 4')
    p = pack_quant_meta(quant_meta, device)
    R, C = quant_meta['qweight'].shape
    out = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in p.items()}
    uniform = int(bits.max()) == int(bits.min())
    out.update({'bits': int(bits.max()) if uniform else None, 'shape': (int(R), int(C))})
    if not uniform:
        out['group_bit_widths'] = bits.detach().to('cpu', torch.uint8).clone()
    return out


def unpack_layer(layer: dict, device: torch.device = torch.device('cuda')) -> dict:
    """Packed layer -> the reference's quant_meta (same 9 keys, dtypes and values as gptq_quant returns), on the CPU."""
    R, C = layer['shape']
    gs = int(layer['group_size'])
    G = C // gs
    codes = _ops.unpack_int4(layer['qweight_packed'].to(device).contiguous()).cpu()
    if layer.get('perm') is not None:
        perm_inv = layer['perm'].to(torch.int64).argsort().to(torch.int16)
    else:
        perm_inv = torch.arange(C, dtype=torch.int16)
    return {
        'qweight': codes,
        'scale': layer['scale_t'].t().contiguous(),
        'qzero': layer['qzero_t'].t().contiguous(),
        'qscale': None,
        'sscale': None,
        'perm_inv': perm_inv,
        'group_sizes': torch.full([G], gs, dtype=torch.int16),
        'group_bit_widths': (layer['group_bit_widths'].to(torch.uint8).clone() if layer.get('group_bit_widths') is not None
                             else torch.full([G], int(layer.get('bits') or 4), dtype=torch.uint8)),
        'scale_bit_width': None,
    }


def pack_results(results: dict, device: torch.device = torch.device('cuda')) -> dict:
    """The reference's results dict (llama.py:82-85) -> packed model dict."""
    return {'format': FORMAT,
            'layers': {name: pack_layer(meta, device) for name, meta in results['data'].items()},
            'metrics': dict(results.get('metrics', {}))}


def save_packed(packed: dict, path: str) -> None:
    if packed.get('format') != FORMAT:
        raise ValueError('not a packed model dict (use pack_results first)')
    torch.save(packed, path)


def load_packed(path: str) -> dict:
    packed = torch.load(path, map_location='cpu', weights_only=False)
    if not isinstance(packed, dict) or packed.get('format') != FORMAT:
        raise ValueError(f'{path}: not a {FORMAT} file')
    return packed


def packed_bytes(packed: dict) -> int:
    n = 0
    for layer in packed['layers'].values():
        n += sum(v.numel() * v.element_size() for v in layer.values() if torch.is_tensor(v))
    return n


def quantized_linears(packed: dict, device: torch.device, dtype: Optional[torch.dtype] = None) -> dict:
    """{name: W4A16Linear} with the packed buffers on `device` (scales cast to `dtype` when given)."""
    out = {}
    for name, layer in packed['layers'].items():
        p = {'qweight_packed': layer['qweight_packed'].to(device),
             'scale_t': layer['scale_t'].to(device=device, dtype=dtype or layer['scale_t'].dtype),
             'qzero_t': layer['qzero_t'].to(device),
             'perm': layer['perm'].to(device) if layer.get('perm') is not None else None,
             'group_size': int(layer['group_size'])}
        out[name] = W4A16Linear(p)
    return out


def apply_to_model(model: nn.Module, packed: dict, device: torch.device, dtype: Optional[torch.dtype] = None) -> nn.Module:
    """Replace every quantised nn.Linear of `model` (names as in results['data'], e.g. 'model.layers.0.mlp.up_proj') by
    its W4A16Linear.  The dense weights are dropped; nothing else of the model is touched."""
    for name, lin in quantized_linears(packed, device, dtype).items():
        parent_name, _, child = name.rpartition('.')
        parent = model.get_submodule(parent_name) if parent_name else model
        old = getattr(parent, child)
        if getattr(old, 'bias', None) is not None:
            raise NotImplementedError(f'{name}: bias is not carried by quant_meta')
        setattr(parent, child, lin)
    return model
