"""B200-native drop-in for the reference's gptq_py.py (HessianHook, gptq_quant).

Same class/function names, fields, arguments, return dict and error behaviour as
/root/reference/gptq_py.py:9-251, so that llama.py / model_utils.py drive it unchanged;
every arithmetic step is a kernel of libgptq_b200.so:

  HessianHook.add_batch   -> K1 tcgen05 SYRK                     (ref gptq_py.py:19-36)
  HessianHook.invert      -> K2 gather/damp + Cholesky + inverse (ref :39-66)
  gptq_quant              -> K3 block solver + K4 trailing GEMM  (ref :125-193)
                             K5 transpose (+ int4 pack)          (ref :216-229)

Host-side torch ops are limited to what the reference also leaves to torch and that is
O(D) or a single elementwise pass (argsort of the diagonal, metrics).
Not implemented (outside SURVEY section 8 scope, never exercised by llama.py): quant_mse,
quant_symmetric, scale_bit_width (EXL2), non-uniform group sizes; they raise
NotImplementedError instead of silently doing something else.
"""
from __future__ import annotations

import torch

import gptq as _ops
from gptq import accumulate_hessian as accumulate_hessian_c, gptq_quantize_range  # noqa: F401  (reference names)
from gptq_triton.accumulate_hessian import accumulate_hessian as accumulate_hessian_t  # noqa: F401


class HessianHook:
    def __init__(self):
        super().__init__()
        self.hessian: torch.Tensor | None = None
        self.n_samples: int = 0
        self.hessian_inv: torch.Tensor | None = None
        self.perm: torch.Tensor | None = None
        self.perm_inv: torch.Tensor | None = None
        self.invert_tries: int = 0

    @torch.no_grad()
    def add_batch(self, inp: torch.Tensor, use_kernel: bool = True) -> None:
        """inp: (..., N, D).  n_samples counts leading-dimension entries (sequences), as the
        reference does (gptq_py.py:25-27)."""
        assert self.hessian_inv is None
        assert not hasattr(inp, 'fake_mode')
        if inp.dim() <= 2:
            inp = inp[None]
        self.n_samples += len(inp)
        if self.hessian is None:
            self.hessian = torch.zeros(inp.size(-1), inp.size(-1), dtype=torch.float32, device=inp.device)
        inp = inp.flatten(end_dim=-2)
        if not use_kernel:
            # the reference's explicit fp32 option (gptq_py.py:31-33), kept for API parity
            inp = inp.to(dtype=torch.float32)
            torch.addmm(self.hessian, inp.t(), inp, beta=1, alpha=1, out=self.hessian)
        else:
            _ops.hessian_syrk(self.hessian, inp.contiguous(), mode=0)   # upper triangle only: invert() reads nothing else

    @torch.no_grad()
    def invert(self, damp_ratio: float = 1e-2, act_order: bool = True) -> torch.Tensor:
        H = self.hessian
        scale = 2.0 / self.n_samples                                   # ref :41
        diag = H.diagonal() * scale
        diag = torch.where(diag == 0.0, torch.ones_like(diag), diag)   # ref :43-44
        if act_order:
            self.perm = torch.argsort(diag, descending=True, stable=True)   # ref :47 (ties: stable here)
            self.perm_inv = torch.argsort(self.perm)                        # ref :49
        damp = float(damp_ratio * diag.mean())                          # ref :52 (one host sync)
        max_try = 100
        tries = 0
        while (max_try := max_try - 1) >= 0:
            tries += 1
            # cumulative damping: try k factors H + k*damp*I (ref :58)
            Hp, U, info = _ops.hessian_invert(H, self.perm, scale, damp * tries)
            if int(info.item()) == 0 and bool(torch.isfinite(U.diagonal()).all()):
                break
        assert max_try >= 0
        self.invert_tries = tries
        self.hessian = Hp                    # scaled, permuted, damped -- what the reference leaves behind
        self.hessian_inv = U
        return self.hessian_inv


def _uniform(t: torch.Tensor, what: str) -> int:
    v = t.tolist()
    if len(set(v)) != 1:
        raise NotImplementedError(f'{what} must be uniform across groups in gptq_b200 (got {sorted(set(v))})')
    return int(v[0])


@torch.no_grad()
def gptq_quant(
        weight: torch.Tensor,
        hessian_hook: HessianHook | None,
        group_sizes: torch.Tensor,
        group_bit_widths: torch.Tensor,
        scale_bit_width: torch.Tensor = None,
        gptq_use_kernel: bool = True,
        gptq_block_sizes: torch.Tensor = None,
        quant_symmetric: bool = False,
        quant_mse: bool = True,
        quant_max_shrink: float = .8,
        quant_n_grid: int = 100,
        quant_norm: float = 2.4,
        quant_use_kernel: bool = False,
        save_device: torch.device = torch.device('cpu'),
) -> dict[str, dict]:
    """gptq_use_kernel=True  -> the reference's kernel-path semantics: static per-group scale/zero
                                from the un-compensated weights, fp16-rounded dequant (ref :125-157)
       gptq_use_kernel=False -> the reference's fp32 PyTorch semantics: per-group scale/zero from
                                the error-compensated weights, fp32 dequant (ref :158-193)
    Both run on the same CUDA block solver."""
    if quant_mse:
        raise NotImplementedError('quant_mse=True (MSE grid search, quant.py:120-165) is outside this build\'s scope')
    if quant_symmetric or (scale_bit_width is not None):
        raise NotImplementedError('symmetric / EXL2 scale quantisation is outside this build\'s scope')
    n_rows, n_columns = weight.shape
    dtype = weight.dtype
    device = weight.device
    gs = _uniform(group_sizes, 'group_sizes')
    bits = _uniform(group_bit_widths, 'group_bit_widths')
    if gptq_block_sizes is not None and _uniform(gptq_block_sizes, 'gptq_block_sizes') != gs:
        raise NotImplementedError('gptq_block_sizes must equal group_sizes (as llama.py:155 sets them)')
    n_groups = n_columns // gs
    assert n_groups * gs == n_columns and len(group_sizes) == n_groups
    maxq = float(2 ** bits - 1)
    perm = hessian_hook.perm if hessian_hook is not None else None
    weight_c = weight.contiguous()

    if hessian_hook is None:
        # round-to-nearest (ref :101-124): the block solver with U = I propagates nothing
        U = torch.eye(n_columns, dtype=torch.float32, device=device)
    else:
        U = hessian_hook.hessian_inv.contiguous()

    # (C, R) fp32, permuted (ref :90-94,128); orig_t / colsq feed the result metrics after the solver has consumed Wt
    Wt, orig_t, colsq = _ops.prep_weight_t2(weight_c, perm)
    if gptq_use_kernel and hessian_hook is not None:
        # static parameters from the original permuted weights (ref :131-134, quant.py:114-118)
        w3 = Wt.view(n_groups, gs, n_rows)
        xmax = w3.amax(dim=1).relu()
        xmin = -(-w3.amin(dim=1)).relu()
        scale_gr = (xmax - xmin) / maxq + 1e-12                               # (G, R)
        qzero_gr = (-xmin / scale_gr).round()
        if gs in (32, 64, 128):
            # the reference's per-group gptq_quantize_range loop (ref :135-139), host loop inside the library
            quant_t, q_t, err_t = _ops.gptq_solve_static(Wt, U, scale_gr.contiguous(), qzero_gr.contiguous(), gs, maxq)
        else:
            quant_t = torch.empty_like(Wt)
            err_t = torch.empty_like(Wt)
            q_t = torch.empty_like(Wt, dtype=torch.int16)
            for g in range(n_groups):
                gptq_quantize_range(quant_t, scale_gr[g].contiguous(), q_t, qzero_gr[g].contiguous(), maxq, U, Wt, err_t,
                                    g * gs, (g + 1) * gs)
        codes_t = q_t.to(torch.uint8)
    else:
        quant_t, codes_t, err_t, scale_gr, qzero_gr = _ops.gptq_solve(Wt, U, gs, maxq)

    # ---- metrics (ref :196-199, :231-237) in one pass over the solver outputs and ONE host sync:
    #   gptq_error = mean(err^2);  gptq_norm = mean((W / diag(U))^2) with W unpermuted against the permuted-order diag(U),
    #   reproduced as is (ref :198) = sum_c colsq_orig[c] / diag(U)[c]^2 / (R C);  l2.0_error = mean((quant - W)^2), which
    #   does not depend on the column order;  l2.0_norm in the weight dtype like the reference.
    sums = _ops.solve_metrics(err_t, quant_t, orig_t)                            # float64 [sum err^2, sum (quant - W)^2]
    if perm is not None:
        colsq_orig = torch.empty_like(colsq)
        colsq_orig[perm] = colsq
    else:
        colsq_orig = colsq
    norm_sum = (colsq_orig.double() / U.diagonal().double() ** 2).sum()
    l2_norm = (weight.abs() ** 2.).mean()                                        # ref :237 (fp16 / bf16 arithmetic)
    n_el = float(n_rows * n_columns)
    e2, d2, ns, l2n = torch.stack([sums[0], sums[1], norm_sum, l2_norm.double()]).tolist()
    if hessian_hook is None:
        gptq_error, gptq_norm = 0.0, 0.0                                         # ref :198 (error == 0)
    else:
        gptq_error, gptq_norm = e2 / n_el, ns / n_el
    metrics: dict[str, float] = {'gptq_error': gptq_error, 'gptq_norm': gptq_norm}

    qweight, _ = _ops.transpose_pack_int4(codes_t, want_codes=True, want_packed=False)   # (R, C) uint8, permuted order
    scale = scale_gr.t().to(dtype=dtype).contiguous()                           # (R, G)
    qzero = qzero_gr.t().to(dtype=torch.uint8).contiguous()                     # (R, G)
    if perm is not None:
        perm16 = perm.to(dtype=torch.int16)
    else:
        perm16 = torch.arange(n_columns, dtype=torch.int16, device=device)
    perm_inv = perm16.argsort().to(dtype=torch.int16)                           # ref :226

    metrics.update({'l2.0_error': d2 / n_el, 'l2.0_norm': l2n})

    quant_meta: dict[str, torch.Tensor] = {
        'qweight': qweight.to(device=save_device),
        'scale': scale.to(device=save_device),
        'qzero': qzero.to(device=save_device),
        'qscale': None,
        'sscale': None,
        'perm_inv': perm_inv.to(device=save_device),
        'group_sizes': group_sizes.to(dtype=torch.int16, device=save_device),
        'group_bit_widths': group_bit_widths.to(dtype=torch.uint8, device=save_device),
        'scale_bit_width': None,
    }
    return {'quant_meta': quant_meta, 'metrics': metrics}
