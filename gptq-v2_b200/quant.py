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

from typing import Optional

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


def quantize2(x, scale, maxq):
    """Quadratic grid for a column of scales (EXL2): clamp(round(sqrt(x / scale)), 1, maxq + 1)   (reference quant.py:40-53)"""
    return torch.clamp(torch.round(torch.sqrt(x / scale)), 1., maxq + 1.)


def dequantize2(qx, scale):
    """qx^2 * scale   (reference quant.py:56-65)"""
    return qx * qx * scale


def dequantize2_quantized2(x, scale, maxq):
    return dequantize2(quantize2(x, scale, maxq), scale)


def exl2_scale_grid(scale, smaxq):
    """The EXL2 re-quantisation of a group's scales (reference quant.py:109-112).  scale (..., R, 1) fp32.
    Returns qscale (..., R, 1), sscale (..., 1, 1), dequantised scale."""
    sscale = scale.amax(dim=-2, keepdim=True) / (smaxq + 1.) ** 2. + EPSILON
    qscale = quantize2(scale, sscale, smaxq)
    return qscale, sscale, dequantize2(qscale, sscale)


def exl2_scale_mse(scale, smaxq, p, norm):
    """Second half of Quantizer.mse (reference quant.py:157-165): one shrink factor for the group's sscale.
    scale (..., R, 1) fp32 (already MSE-refined), p (Q) shrink factors."""
    sscale = scale.amax(dim=-2, keepdim=True) / (smaxq + 1.) ** 2. + EPSILON
    cand = sscale[..., None] * p
    q = dequantize2(quantize2(scale[..., None], cand, smaxq), cand)                                    # (..., R, 1, Q)
    arg = (q - scale[..., None]).abs().pow(norm).sum(dim=-3).argmin(dim=-1, keepdim=True)              # (..., 1, 1)
    sscale = sscale * p[arg]
    qscale = quantize2(scale, sscale, smaxq)
    return qscale, sscale, dequantize2(qscale, sscale)


class Quantizer:
    """Per-row quantiser over the last dimension with the reference's options (quant.py:72-183): asymmetric min/max or
    symmetric grid, EXL2 quantised scales (qscale / sscale), MSE refinement of the scale.  Plain torch on whatever device
    x lives on: this class is the host mirror for callers of the reference API; gptq_quant itself runs the same
    arithmetic in libgptq_b200.so (gq_find_params / gq_mse_refine) on the solver's transposed layout."""

    def __init__(self, scale=None, qzero=None, maxq=None, qscale=None, sscale=None, smaxq=None):
        self.scale, self.qzero, self.maxq = scale, qzero, maxq
        self.qscale, self.sscale, self.smaxq = qscale, sscale, smaxq
        if self.scale is None and self.qscale is not None and self.sscale is not None:
            self.scale = dequantize2(self.qscale, self.sscale)
        elif self.qscale is None and self.scale is not None and self.sscale is not None and self.smaxq is not None:
            self.qscale = quantize2(self.scale, self.sscale, self.smaxq)

    def find_params(self, x, bit_width, sym: bool = False, scale_bit_width=None) -> None:
        exl2 = scale_bit_width is not None and scale_bit_width > 0
        self.maxq = 2. ** bit_width - 1.
        if sym or exl2:
            self.scale = x.abs().amax(dim=-1, keepdim=True) * (2. / self.maxq) + EPSILON
            mid = (self.maxq + 1.) * .5
            self.qzero = torch.full_like(self.scale, float(torch.round(torch.as_tensor(mid))))
            if exl2:
                self.smaxq = 2. ** scale_bit_width - 1.
                self.qscale, self.sscale, self.scale = exl2_scale_grid(self.scale, self.smaxq)
        else:
            hi = torch.relu(x.amax(dim=-1, keepdim=True))
            lo = -torch.relu(-x.amin(dim=-1, keepdim=True))
            self.scale = (hi - lo) / self.maxq + EPSILON
            self.qzero = torch.round(-lo / self.scale)

    def mse(self, x, max_shrink: float = .8, n_grid: int = 100, norm: float = 2.4, use_kernel: bool = False) -> None:
        if use_kernel:
            # the reference calls exllamav2's ext_c.quantize_err here, whose import is commented out (quant.py:4)
            raise NotImplementedError('Quantizer.mse(use_kernel=True) needs exllamav2, which the reference itself does not import')
        p = 1. - torch.arange(0., max_shrink, 1. / n_grid, device=x.device)
        q = dequantize_quantized(x[..., None], self.scale[..., None] * p, self.qzero[..., None], self.maxq)
        best = (q - x[..., None]).abs().pow(norm).sum(dim=-2).argmin(dim=-1, keepdim=True)
        self.scale = self.scale * p[best]
        if self.sscale is not None:
            self.qscale, self.sscale, self.scale = exl2_scale_mse(self.scale, self.smaxq, p, norm)

    def quantize(self, x):
        return quantize(x, self.scale, self.qzero, self.maxq)

    def dequantize(self, qx):
        return dequantize(qx, self.scale, self.qzero)

    def dequantize_quantized(self, x):
        return dequantize_quantized(x, self.scale, self.qzero, self.maxq)


def _uniform_group(group_sizes: torch.Tensor) -> int:
    v = set(group_sizes.tolist())
    if len(v) != 1:
        raise NotImplementedError('the packed int4 format / K6 take one group size per linear (llama.py:146 uses 128 everywhere); '
                                  'reconstruct_nn_linear handles non-uniform groups densely')
    return int(v.pop())


def _expand_groups(per_group: torch.Tensor, group_sizes: torch.Tensor) -> torch.Tensor:
    """(R, G) per-group values -> (R, C) per-column values for any group sizes (the reference loops over the groups,
    quant.py:230-236)."""
    sizes = group_sizes.to(device=per_group.device, dtype=torch.int64)
    if int(sizes.min()) == int(sizes.max()):
        gs = int(sizes[0])
        return per_group.unsqueeze(-1).expand(*per_group.shape, gs).reshape(per_group.shape[0], -1)
    return per_group.repeat_interleave(sizes, dim=1)


def construct_matrix(qweight, scale, qzero, group_sizes):
    """Dense (R, C) matrix in scale.dtype from codes (permuted column order kept).
    int16 subtract, then multiply in the 16-bit dtype, as reference quant.py:222-237.  Any group sizes."""
    q = qweight.to(torch.int16)
    z = _expand_groups(qzero.to(torch.int16), group_sizes)
    return (q - z) * _expand_groups(scale, group_sizes)


def construct_matrix_2(qweight, qzero, qscale, sscale, group_sizes):
    """EXL2 metadata (reference quant.py:240-259): scale = qscale^2 (int16) * sscale, evaluated in sscale.dtype."""
    qs = qscale.to(torch.int16)
    scale = qs * qs * sscale.reshape(1, -1)
    q = qweight.to(torch.int16)
    z = _expand_groups(qzero.to(torch.int16), group_sizes)
    return (q - z) * _expand_groups(scale, group_sizes)


def _fused_reconstruction(quant_meta: dict, dtype: torch.dtype, device: torch.device, group_size: Optional[int]):
    """The dense weight through gq_reconstruct_dense (one launch, no host synchronisation) when the metadata allows it:
    CUDA target, 16-bit dtype, plain scales, and a uniform group size that is known WITHOUT reading a device tensor (the
    caller's `group_size`, or group_sizes held on the host).  None otherwise: the torch route below handles the rest."""
    if torch.device(device).type != 'cuda' or dtype not in (torch.float16, torch.bfloat16) or quant_meta.get('sscale') is not None:
        return None
    qweight = quant_meta['qweight']
    R, C = qweight.shape
    if group_size is None:
        gs = quant_meta['group_sizes']
        if gs.is_cuda or gs.numel() == 0 or int(gs.min()) != int(gs.max()):
            return None
        group_size = int(gs[0])
    group_size = int(group_size)
    if C % 16 or C > 32767 or group_size < 1 or C % group_size or C // group_size != quant_meta['scale'].shape[1]:
        return None
    perm_inv = quant_meta.get('perm_inv')
    if perm_inv is not None:
        perm_inv = perm_inv.to(device=device, dtype=torch.int16).contiguous()
    return _ops.reconstruct_dense(qweight.to(device=device).contiguous(), quant_meta['scale'].to(dtype=dtype, device=device).contiguous(),
                                  quant_meta['qzero'].to(device=device).contiguous(), perm_inv, group_size)


def reconstruct_nn_linear(quant_meta: dict, dtype: torch.dtype, device: torch.device = torch.device('cpu'), *,
                          group_size: Optional[int] = None) -> nn.Linear:
    """Reference-compatible dense reconstruction (quant.py:262-279).  `group_size` (keyword-only, not in the reference): the
    caller's promise that all groups have this size, which lets a CUDA reconstruction run as one kernel without reading
    quant_meta['group_sizes'] back from the device (llama_driver.py passes it)."""
    fused = _fused_reconstruction(quant_meta, dtype, device, group_size)
    if fused is not None:
        lin = nn.Linear(fused.shape[1], fused.shape[0], bias=False, dtype=dtype, device='meta')
        lin.weight = nn.Parameter(fused)                          # requires_grad as nn.Linear's default, like the reference's
        lin.eval()
        return lin
    qweight = quant_meta['qweight'].to(device=device)
    if quant_meta.get('sscale') is not None:
        weight = construct_matrix_2(qweight, quant_meta['qzero'].to(device=device), quant_meta['qscale'].to(device=device),
                                    quant_meta['sscale'].to(dtype=dtype, device=device), quant_meta['group_sizes'])
    else:
        weight = construct_matrix(qweight, quant_meta['scale'].to(dtype=dtype, device=device),
                                  quant_meta['qzero'].to(device=device), quant_meta['group_sizes'])
    if quant_meta.get('perm_inv') is not None:
        weight = weight[:, quant_meta['perm_inv'].to(dtype=torch.int32, device=device)]
    lin = nn.Linear(weight.shape[1], weight.shape[0], bias=False, dtype=dtype, device=weight.device)
    lin.weight.data = weight.contiguous()
    lin.eval()
    return lin


def _meta_scale(quant_meta: dict, device: torch.device) -> torch.Tensor:
    """(R, G) 16-bit scale a dense reconstruction would use: quant_meta['scale'], or qscale^2 * sscale for EXL2 metadata."""
    if quant_meta.get('sscale') is None:
        return quant_meta['scale'].to(device=device)
    qs = quant_meta['qscale'].to(device=device, dtype=torch.int16)
    return qs * qs * quant_meta['sscale'].to(device=device).reshape(1, -1)


def pack_quant_meta(quant_meta: dict, device: torch.device) -> dict:
    """quant_meta (one uint8 per code) -> packed storage for K6:
    qweight_packed (R, C/8) int32, scale_t (G, R) and qzero_t (G, R) uint8 (group-major), perm (C) int32."""
    gs = _uniform_group(quant_meta['group_sizes'])
    if int(quant_meta['group_bit_widths'].max()) > 4:
        raise NotImplementedError('packed storage / K6 hold one code per nibble: bit widths up to 4 (llama.py:147 uses 4)')
    codes = quant_meta['qweight'].to(device=device).contiguous()
    perm = None
    if quant_meta.get('perm_inv') is not None:
        perm = quant_meta['perm_inv'].to(device=device, dtype=torch.int64).argsort().to(torch.int32).contiguous()
    return {
        'qweight_packed': _ops.pack_int4(codes),
        'scale_t': _meta_scale(quant_meta, device).t().contiguous(),
        'qzero_t': quant_meta['qzero'].to(device=device).t().contiguous(),
        'perm': perm,
        'group_size': gs,
    }


class W4A16Linear(nn.Module):
    """y = x W_hat^T from packed int4 codes.  Batches below `dense_rows(N, K)` rows run K6 (dequantisation fused into the
    GEMM, no dense copy of W_hat exists).  Larger batches are compute-bound and K6 would dequantise the weights again
    for every 256 batch columns: there the packed weights are dequantised ONCE into a transient 16-bit matrix
    (gq_w4a16_dequant, the reference's construct_matrix rounding) and multiplied by this build's dense tcgen05 GEMM
    (gq_gemm_f16_nt) -- what the reference does for every batch size with cuBLAS (quant.py:262-279 + nn.Linear), but from
    4x fewer resident bytes and without a library call.  Crossovers measured on
    B200 (profiles/r01_v29_w4a16_sweep.jsonl): 128 rows for N*K <= 2^26 (all Llama-2-7B shapes), 512 rows above."""
    dense_route = True

    @staticmethod
    def dense_rows(N: int, K: int) -> int:
        return 128 if N * K <= (1 << 26) else 512

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
        if self.dense_route and x2.shape[0] >= self.dense_rows(self.out_features, self.in_features):
            dense = _ops.w4a16_dequant(self.qweight_packed, scale_t, self.qzero_t, self.group_size)   # transient
            xp = x2.index_select(1, self.perm) if self.perm is not None else x2
            y = _ops.gemm_f16_nt(xp.contiguous(), dense)                                              # own tcgen05 GEMM, no library call
        else:
            y = _ops.w4a16_gemm(x2, self.qweight_packed, scale_t, self.qzero_t, self.perm, self.group_size)
        return y.view(*lead, self.out_features)
