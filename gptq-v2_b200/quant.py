"""Host-side mirror of the parts of the reference's quant.py that sit on the GPTQ path
(/root/reference/quant.py:10-30, 89-118, 222-237, 262-279), plus what the reference lacks:
packed int4 storage and a module that evaluates a quantised linear through the W4A16
kernel instead of a dense fp16 reconstruction.

`reconstruct_nn_linear` keeps the reference signature and result (a dense nn.Linear in the
model dtype) because llama.py:172-177 replays calibration batches through it; the dense
matrix is produced by unpack -> elementwise torch ops exactly as the reference does.
`W4A16Linear.from_quant_meta` is the drop-in for serving the quantised model.
"""
from __future__ import annotations

import torch
from torch import nn

import gptq as _ops

EPSILON = 1e-12


def quantize(x, scale, qzero, maxq):
    """codes = clamp(round(x / scale) + qzero, 0, maxq)   (reference quant.py:19)"""
    return torch.clamp(torch.round(x / scale) + qzero, 0., maxq)


def dequantize(qx, scale, qzero):
    """(codes - qzero) * scale   (reference quant.py:30)"""
    return (qx - qzero) * scale


def dequantize_quantized(x, scale, qzero, maxq):
    return dequantize(quantize(x, scale, qzero, maxq), scale, qzero)


class Quantizer:
    """Asymmetric min/max quantiser over the last dimension (reference quant.py:72-118).
    Symmetric / EXL2 / MSE variants are outside this build's scope and raise."""

    def __init__(self, scale=None, qzero=None, maxq=None):
        self.scale, self.qzero, self.maxq = scale, qzero, maxq

    def find_params(self, x, bit_width, sym: bool = False, scale_bit_width=None) -> None:
        if sym or (scale_bit_width is not None and scale_bit_width > 0):
            raise NotImplementedError('symmetric / EXL2 quantisation is outside gptq_b200\'s scope')
        self.maxq = 2. ** bit_width - 1.
        hi = torch.relu(x.amax(dim=-1, keepdim=True))
        lo = -torch.relu(-x.amin(dim=-1, keepdim=True))
        self.scale = (hi - lo) / self.maxq + EPSILON
        self.qzero = torch.round(-lo / self.scale)

    def mse(self, *args, **kwargs) -> None:
        raise NotImplementedError('MSE grid search is outside gptq_b200\'s scope')

    def quantize(self, x):
        return quantize(x, self.scale, self.qzero, self.maxq)

    def dequantize(self, qx):
        return dequantize(qx, self.scale, self.qzero)

    def dequantize_quantized(self, x):
        return dequantize_quantized(x, self.scale, self.qzero, self.maxq)


def _uniform_group(group_sizes: torch.Tensor) -> int:
    v = set(group_sizes.tolist())
    if len(v) != 1:
        raise NotImplementedError('non-uniform group sizes are outside gptq_b200\'s scope')
    return int(v.pop())


def construct_matrix(qweight, scale, qzero, group_sizes):
    """Dense (R, C) matrix in scale.dtype from codes (permuted column order kept).
    int16 subtract, then multiply in the 16-bit dtype, as reference quant.py:222-237."""
    gs = _uniform_group(group_sizes)
    R, C = qweight.shape
    q = qweight.to(torch.int16).view(R, C // gs, gs)
    z = qzero.to(torch.int16).expand(scale.shape).unsqueeze(-1)
    return ((q - z) * scale.unsqueeze(-1)).view(R, C)


def reconstruct_nn_linear(quant_meta: dict, dtype: torch.dtype, device: torch.device = torch.device('cpu')) -> nn.Linear:
    """Reference-compatible dense reconstruction (quant.py:262-279)."""
    if quant_meta.get('sscale') is not None:
        raise NotImplementedError('EXL2 metadata is outside gptq_b200\'s scope')
    qweight = quant_meta['qweight'].to(device=device)
    weight = construct_matrix(qweight, quant_meta['scale'].to(dtype=dtype, device=device),
                              quant_meta['qzero'].to(device=device), quant_meta['group_sizes'])
    if quant_meta.get('perm_inv') is not None:
        weight = weight[:, quant_meta['perm_inv'].to(dtype=torch.int32, device=device)]
    lin = nn.Linear(weight.shape[1], weight.shape[0], bias=False, dtype=dtype, device=weight.device)
    lin.weight.data = weight.contiguous()
    lin.eval()
    return lin


def pack_quant_meta(quant_meta: dict, device: torch.device) -> dict:
    """quant_meta (one uint8 per code) -> packed storage for K6:
    qweight_packed (R, C/8) int32, scale_t (G, R) and qzero_t (G, R) uint8 (group-major), perm (C) int32."""
    gs = _uniform_group(quant_meta['group_sizes'])
    codes = quant_meta['qweight'].to(device=device).contiguous()
    perm = None
    if quant_meta.get('perm_inv') is not None:
        perm = quant_meta['perm_inv'].to(device=device, dtype=torch.int64).argsort().to(torch.int32).contiguous()
    return {
        'qweight_packed': _ops.pack_int4(codes),
        'scale_t': quant_meta['scale'].to(device=device).t().contiguous(),
        'qzero_t': quant_meta['qzero'].to(device=device).t().contiguous(),
        'perm': perm,
        'group_size': gs,
    }


class W4A16Linear(nn.Module):
    """y = x W_hat^T straight from packed int4 codes (K6); no dense copy of W_hat exists."""

    def __init__(self, packed: dict):
        super().__init__()
        self.register_buffer('qweight_packed', packed['qweight_packed'])
        self.register_buffer('scale_t', packed['scale_t'])
        self.register_buffer('qzero_t', packed['qzero_t'])
        self.register_buffer('perm', packed['perm'])
        self.group_size = packed['group_size']
        self.out_features, self.in_features = self.qweight_packed.shape[0], self.qweight_packed.shape[1] * 8

    @classmethod
    def from_quant_meta(cls, quant_meta: dict, device: torch.device) -> 'W4A16Linear':
        return cls(pack_quant_meta(quant_meta, device))

    @torch.no_grad()
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self.in_features).contiguous()
        scale_t = self.scale_t if self.scale_t.dtype == x2.dtype else self.scale_t.to(x2.dtype)
        y = _ops.w4a16_gemm(x2, self.qweight_packed, scale_t, self.qzero_t, self.perm, self.group_size)
        return y.view(*lead, self.out_features)
