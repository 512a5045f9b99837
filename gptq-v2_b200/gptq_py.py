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
The Quantizer options llama.py never sets are supported on all three branches:
quant_symmetric, quant_mse (grid search, gq_mse_refine), scale_bit_width (EXL2 qscale / sscale), per-group bit widths.
Group sizes need not be uniform (any sizes on the round-to-nearest branch, multiples of 32 on the two GPTQ branches), and
on the fp32 branch gptq_block_sizes may differ from group_sizes in any way (nested either way or not at all, multiples of
32), with every Quantizer option: the MSE search then runs over the block's columns as in ref :178-184.
Still NotImplementedError: quant_use_kernel=True (needs exllamav2, which the reference itself no longer imports).
"""
from __future__ import annotations

import collections.abc

import torch

import gptq as _ops
from gptq import accumulate_hessian as accumulate_hessian_c, gptq_quantize_range  # noqa: F401  (reference names)
from gptq_triton.accumulate_hessian import accumulate_hessian as accumulate_hessian_t  # noqa: F401
from quant import exl2_scale_grid, exl2_scale_mse


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
    def invert(self, damp_ratio: float = 1e-2, act_order: bool = True, *, defer_check: bool = False) -> torch.Tensor:
        """ref :39-66.  defer_check (not in the reference): queue the first factorisation and return without the host
        synchronisation that decides about a damping retry; the caller queues the work that consumes `hessian_inv` behind it
        and then calls settle(), which does the check (and the reference's retry loop if the first try failed)."""
        H = self.hessian
        scale = 2.0 / self.n_samples                                   # ref :41
        diag = H.diagonal() * scale
        diag = torch.where(diag == 0.0, torch.ones_like(diag), diag)   # ref :43-44
        if act_order:
            self.perm = torch.argsort(diag, descending=True, stable=True)   # ref :47 (ties: stable here)
            # ref :49 takes argsort(perm); the inverse of a permutation is one scatter
            self.perm_inv = torch.empty_like(self.perm)
            self.perm_inv[self.perm] = torch.arange(self.perm.numel(), device=self.perm.device, dtype=self.perm.dtype)
        damp = (damp_ratio * diag.mean()).reshape(())                   # ref :52; stays on the device: K2 reads it there
        self._pending = (H, scale, damp, 1, None)
        self._factor_once()
        if not defer_check:
            self.settle()
        return self.hessian_inv

    def _factor_once(self) -> None:
        H, scale, damp, tries, _ = self._pending
        # cumulative damping: try k factors H + k*damp*I (ref :58)
        Hp, U, info = _ops.hessian_invert(H, self.perm, scale, damp, damp_mult=float(tries))
        self._pending = (H, scale, damp, tries, info)
        self.invert_tries = tries
        self.hessian = Hp                    # scaled, permuted, damped -- what the reference leaves behind
        self.hessian_inv = U

    def settle(self) -> bool:
        """The check invert(defer_check=True) left out: one host synchronisation; on a non-positive pivot or a non-finite
        diag(U) the reference's retry loop (ref :55-65) runs now.  Returns True when the FIRST factorisation was good, i.e.
        when everything that was queued behind it used the final `hessian_inv`; False means it has to be redone."""
        pending = getattr(self, '_pending', None)
        if pending is None:
            return True
        first_ok = True
        max_try = 100
        while True:
            H, scale, damp, tries, info = self._pending
            if not bool((info != 0) | ~torch.isfinite(self.hessian_inv.diagonal()).all()):   # one host sync per try (ref :63)
                break
            first_ok = False
            max_try -= 1
            assert max_try > 0
            self._pending = (H, scale, damp, tries + 1, None)
            self._factor_once()
        self._pending = None
        return first_ok


def _mean_in_dtype(total: float, n: float, dtype: torch.dtype) -> float:
    """mean of n values whose (fp32-accumulated) sum is `total`, rounded to `dtype` like torch's reduction (ref :237 takes
    (weight.abs() ** 2.).mean() in the weight dtype)."""
    return float(torch.tensor(total / n, dtype=torch.float32).to(dtype))


def _metrics_from_sums(sums, n_el: float, has_hook: bool, dtype: torch.dtype) -> dict:
    """[sum err^2, sum (quant - W)^2, gptq_norm sum, sum dtype(w^2)] -> the reference's four metrics (ref :196-199, :231-237)."""
    e2, d2, ns, l2s = sums
    return {'gptq_error': e2 / n_el if has_hook else 0.0, 'gptq_norm': ns / n_el if has_hook else 0.0,      # ref :198 (0 without H)
            'l2.0_error': d2 / n_el, 'l2.0_norm': _mean_in_dtype(l2s, n_el, dtype)}


class DeferredMetrics(collections.abc.Mapping):
    """The four metrics of gptq_quant (ref :196-199, :231-237) as a read-only mapping whose values are fetched from the device
    on first access.  gptq_quant(defer_metrics=True) returns it instead of a dict, so that a caller quantising several linears
    in a row (sharded.gptq_quant_group_sharded, bench.py, llama_driver.py) does not synchronise the host after every linear: the
    four sums travel to pinned host memory behind the solver and are read when somebody looks."""
    KEYS = ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')

    def __init__(self, stats: torch.Tensor, n_el: float, has_hook: bool, dtype: torch.dtype):
        self._stats, self._n_el, self._has_hook, self._dtype = stats, n_el, has_hook, dtype
        self._host = torch.empty(4, dtype=torch.float64, pin_memory=True)
        self._host.copy_(stats, non_blocking=True)
        self._event = torch.cuda.Event()
        self._event.record()
        self._values = None

    def device_values(self) -> torch.Tensor:
        """(4,) float64 on the device, in KEYS order (for reductions that stay on the device)."""
        e2, d2, ns, l2s = self._stats.unbind(0)
        zero = torch.zeros_like(e2)
        l2n = (l2s / self._n_el).to(torch.float32).to(self._dtype).double()
        return torch.stack([e2 / self._n_el if self._has_hook else zero, ns / self._n_el if self._has_hook else zero,
                            d2 / self._n_el, l2n])

    def _resolve(self) -> dict:
        if self._values is None:
            self._event.synchronize()
            self._values = _metrics_from_sums(self._host.tolist(), self._n_el, self._has_hook, self._dtype)
        return self._values

    def __getitem__(self, key):
        return self._resolve()[key]

    def __iter__(self):
        return iter(self.KEYS)

    def __len__(self):
        return len(self.KEYS)

    def __repr__(self):
        return repr(self._resolve())


def _int_list(t) -> list[int]:
    """group_sizes / group_bit_widths / gptq_block_sizes as Python ints.  Pass CPU tensors (or lists): a CUDA tensor costs a
    device synchronisation here, like the reference's own .tolist() (gptq_py.py:98)."""
    return [int(v) for v in (t.tolist() if isinstance(t, torch.Tensor) else t)]


def _runs(sizes: list[int]):
    """Consecutive groups of equal size as (first group, number of groups, group size, first column)."""
    out, g, c = [], 0, 0
    while g < len(sizes):
        n = 1
        while g + n < len(sizes) and sizes[g + n] == sizes[g]:
            n += 1
        out.append((g, n, sizes[g], c))
        c += n * sizes[g]
        g += n
    return out


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
        *,
        defer_metrics: bool = False,
        _row_segments=None,
) -> dict[str, dict]:
    """gptq_use_kernel=True  -> the reference's kernel-path semantics: static per-group scale/zero
                                from the un-compensated weights, fp16-rounded dequant (ref :125-157)
       gptq_use_kernel=False -> the reference's fp32 PyTorch semantics: per-group scale/zero from
                                the error-compensated weights, fp32 dequant (ref :158-193)
       hessian_hook=None     -> round to nearest (ref :101-124)
    Group sizes need not be uniform (ref :98 walks a cumsum); on the fp32 branch GPTQ blocks may differ from the groups
    (ref :161-193).  Group parameters come from the solver itself (asymmetric min/max) or from gq_find_params /
    gq_mse_refine (symmetric, MSE-refined, EXL2 grids).
    defer_metrics (not in the reference): return the metrics as a DeferredMetrics mapping instead of floats, i.e. without the
    host synchronisation at the end of the call."""
    if quant_mse and quant_use_kernel:
        # ref quant.py:131-150 calls exllamav2's ext_c.quantize_err, whose import the reference has commented out (quant.py:4)
        raise NotImplementedError('quant_use_kernel=True needs exllamav2 (not importable in the reference either)')
    n_rows, n_columns = weight.shape
    dtype = weight.dtype
    device = weight.device
    sizes = _int_list(group_sizes)
    bits_list = _int_list(group_bit_widths)
    n_groups = len(sizes)
    offs = [0]
    for n in sizes:
        offs.append(offs[-1] + n)
    assert offs[-1] == n_columns and len(bits_list) == n_groups and min(sizes) > 0
    uniform = len(set(sizes)) == 1
    gs = sizes[0] if uniform else None
    maxq_list = [float(2 ** b - 1) for b in bits_list]
    sbits = int(scale_bit_width) if scale_bit_width is not None else None
    exl2 = sbits is not None and sbits > 0                                  # ref quant.py:99-100,108
    sym = bool(quant_symmetric) or exl2
    plain = not (sym or quant_mse)                                          # llama.py's setting: asymmetric min/max, no refinement
    perm = hessian_hook.perm if hessian_hook is not None else None
    weight_c = weight.contiguous()
    U = hessian_hook.hessian_inv.contiguous() if hessian_hook is not None else None

    # (C, R) fp32, permuted (ref :90-94,128); orig_t / colsq feed the result metrics after the solver has consumed Wt
    Wt, orig_t, _ = _ops.prep_weight_t2(weight_c, perm, getattr(hessian_hook, 'perm_inv', None) if perm is not None else None,
                                        want_colsq=False)
    scale_gr = torch.empty(n_groups, n_rows, dtype=torch.float32, device=device)
    qzero_gr = torch.empty_like(scale_gr)
    qscale_gr = sscale_g = None
    solver_gs = uniform and gs in (32, 64, 128)
    fast = (hessian_hook is not None and not gptq_use_kernel and plain and solver_gs
            and (gptq_block_sizes is None or _int_list(gptq_block_sizes) == sizes))
    if not fast:
        maxq_dev = torch.tensor(maxq_list, dtype=torch.float32, device=device)
        if exl2:
            qscale_gr = torch.empty_like(scale_gr)
            sscale_g = torch.empty(n_groups, dtype=torch.float32, device=device)
        grid = (1. - torch.arange(0., quant_max_shrink, 1. / quant_n_grid, device=device)) if quant_mse else None   # ref quant.py:152
        smaxq = torch.tensor(2. ** sbits - 1., dtype=torch.float32, device=device) if exl2 else None

        def group_params(g0: int, n: int, mse_cols=None) -> None:
            """Quantizer.find_params (+ .mse) of the equal-sized groups [g0, g0 + n) from the current Wt (ref quant.py:89-165):
            the O(R C) and O(R C Q) passes run in the library, the (R)-sized EXL2 scale grid in torch on the device.
            mse_cols = (i1, i2): the MSE search sums the error over those columns (the GPTQ block, ref :178-184) instead
            of the group's own (n == 1 then)."""
            size, c0 = sizes[g0], offs[g0]
            _ops.find_params_cols(Wt, c0, size, n, maxq_dev, g0, sym, scale_gr, qzero_gr)
            rows = slice(g0, g0 + n)
            if exl2:
                qs, ss, sc = exl2_scale_grid(scale_gr[rows].unsqueeze(-1), smaxq)
                scale_gr[rows] = sc.squeeze(-1)
            if quant_mse:
                if mse_cols is None:
                    _ops.mse_refine_cols(Wt, c0, size, n, maxq_dev, g0, grid, quant_norm, scale_gr, qzero_gr)
                else:
                    assert n == 1
                    _ops.mse_refine_cols(Wt, mse_cols[0], mse_cols[1] - mse_cols[0], 1, maxq_dev, g0, grid, quant_norm,
                                         scale_gr, qzero_gr)
                if exl2:
                    qs, ss, sc = exl2_scale_mse(scale_gr[rows].unsqueeze(-1), smaxq, grid, quant_norm)
                    scale_gr[rows] = sc.squeeze(-1)
            if exl2:
                qscale_gr[rows] = qs.squeeze(-1)
                sscale_g[rows] = ss.reshape(-1)

    err_t = None
    if hessian_hook is None:
        # ---- round to nearest (ref :101-124): parameters from the weights as they are, element-wise quantisation; no U,
        # no propagation, any group sizes
        quant_t = torch.empty_like(Wt)
        codes_t = torch.empty(n_columns, n_rows, dtype=torch.uint8, device=device)
        for g0, n, size, c0 in _runs(sizes):
            group_params(g0, n)
            _ops.rtn_quantize(Wt, c0, size, n, maxq_dev, g0, scale_gr, qzero_gr, quant_t, codes_t)
    elif gptq_use_kernel:
        # ---- static parameters from the original permuted weights (ref :131-134), then the reference's per-group
        # gptq_quantize_range loop (ref :135-139): one library call when the groups are uniform
        if any(n % 32 for n in sizes):
            raise NotImplementedError('the kernel branch needs group sizes that are multiples of 32')
        for g0, n, size, c0 in _runs(sizes):
            group_params(g0, n)
        if solver_gs:
            quant_t, q_t, err_t = _ops.gptq_solve_ex(Wt, U, scale_gr, qzero_gr, gs, maxq_list, mode=3)
        else:
            quant_t = torch.empty_like(Wt)
            err_t = torch.empty_like(Wt)
            q_t = torch.empty_like(Wt, dtype=torch.int16)
            for g in range(n_groups):
                gptq_quantize_range(quant_t, scale_gr[g], q_t, qzero_gr[g], maxq_list[g], U, Wt, err_t, offs[g], offs[g + 1])
        codes_t = q_t.to(torch.uint8)
    elif fast:
        # ---- min/max parameters of every group computed by the solver itself on the compensated weights (ref :169-176)
        quant_t, codes_t, err_t = _ops.gptq_solve_ex(Wt, U, scale_gr, qzero_gr, gs, maxq_list, mode=0)
    else:
        # ---- fp32 branch, general form (ref :161-193).  The reference computes a group's parameters when its column loop
        # reaches the group's first column, from the GLOBAL weights: they hold the lazy-batch updates of the finished
        # blocks but not the in-block updates of the current one.  So: at every block start the parameters of all groups
        # that begin inside the block are computed from Wt as it is (the MSE search over the block's columns, ref
        # :178-184), then the block is solved in steps of `step` columns with supplied parameters (mode 4), every step
        # pushing its update to all later columns -- the reference's in-block propagation applied in another order
        # (fp32 reordering only).  Blocks and groups may nest either way or not at all; every boundary must be a
        # multiple of step = gcd(all sizes, 128) in {32, 64, 128}.
        import math
        blocks = _int_list(gptq_block_sizes) if gptq_block_sizes is not None else sizes
        assert sum(blocks) == n_columns and min(blocks) > 0
        step = math.gcd(128, *sizes, *blocks)
        if step not in (32, 64, 128):
            raise NotImplementedError('group sizes and gptq_block_sizes must be multiples of 32 on the fp32 branch')
        n_steps = n_columns // step
        direct = uniform and gs == step                                     # the per-step layout IS the per-group layout
        scale_st = scale_gr if direct else torch.empty(n_steps, n_rows, dtype=torch.float32, device=device)
        qzero_st = qzero_gr if direct else torch.empty_like(scale_st)
        grp_of_step = [g for g in range(n_groups) for _ in range(sizes[g] // step)]
        maxq_st = [maxq_list[g] for g in grp_of_step]
        out, gi, i1 = None, 0, 0
        for bsz in blocks:
            i2 = i1 + bsz
            while gi < n_groups and offs[gi] < i2:
                group_params(gi, 1, mse_cols=(i1, i2) if (quant_mse and (i1, i2) != (offs[gi], offs[gi + 1])) else None)
                if not direct:
                    s0, s1 = offs[gi] // step, offs[gi + 1] // step
                    scale_st[s0:s1] = scale_gr[gi]
                    qzero_st[s0:s1] = qzero_gr[gi]
                gi += 1
            out = _ops.gptq_solve_ex(Wt, U, scale_st, qzero_st, step, maxq_st, mode=4, a=i1, b=i2, out=out)
            i1 = i2
        quant_t, codes_t, err_t = out

    # ---- metrics (ref :196-199, :231-237) in one pass over the solver outputs and ONE host sync:
    #   gptq_error = mean(err^2);  gptq_norm = mean((W / diag(U))^2) with W unpermuted against the permuted-order diag(U),
    #   reproduced as is (ref :198) = sum_c colsq_orig[c] / diag(U)[c]^2 / (R C);  l2.0_error = mean((quant - W)^2), which
    #   does not depend on the column order;  l2.0_norm in the weight dtype like the reference.
    # One library call queues all of it (gq_finalize_quant): the four sums and the (R, G) scale / zero tables of quant_meta.
    seg_rows = [n_rows] if _row_segments is None else [int(n) for n in _row_segments]
    stats, scale, qzero = _ops.finalize_quant(err_t, quant_t, orig_t, scale_gr, qzero_gr, perm, U, seg_rows)   # stats (n_seg, 4)

    qweight, _ = _ops.transpose_pack_int4(codes_t, want_codes=True, want_packed=False)   # (R, C) uint8, permuted order
    qscale = qscale_gr.t().to(dtype=torch.uint8).contiguous() if exl2 else None  # (R, G)   ref :217
    sscale = sscale_g.to(dtype=dtype) if exl2 else None                          # (G)      ref :218
    if perm is not None:
        perm16 = perm.to(dtype=torch.int16)
    else:
        perm16 = torch.arange(n_columns, dtype=torch.int16, device=device)
    bits_out = group_bit_widths.to(dtype=torch.uint8)
    # with EXL2 metadata the groups are stored by descending (bit width, group size), ties in ascending position (ref :201-212)
    order = list(range(n_groups))
    if scale_bit_width is not None:
        order.sort(key=lambda g: (bits_list[g], sizes[g], -g), reverse=True)
    sizes_out = torch.tensor(sizes, dtype=torch.int16)
    if order != list(range(n_groups)):
        ig = torch.tensor(order, dtype=torch.int64, device=device)
        ic = torch.cat([torch.arange(offs[g], offs[g + 1], device=device) for g in order])
        sizes_out = sizes_out[ig.cpu()]
        qweight, perm16 = qweight[:, ic].contiguous(), perm16[ic]
        scale, qzero, bits_out = scale[:, ig].contiguous(), qzero[:, ig].contiguous(), bits_out[ig.to(bits_out.device)]
        if exl2:
            qscale, sscale = qscale[:, ig].contiguous(), sscale[ig]
    if order == list(range(n_groups)) and perm is not None and getattr(hessian_hook, 'perm_inv', None) is not None:
        perm_inv = hessian_hook.perm_inv.to(dtype=torch.int16)                  # == perm16.argsort() (ref :226) without the sort
    else:
        perm_inv = perm16.argsort().to(dtype=torch.int16)                       # ref :226

    quant_meta: dict[str, torch.Tensor] = {
        'qweight': qweight.to(device=save_device),
        'scale': scale.to(device=save_device),
        'qzero': qzero.to(device=save_device),
        'qscale': qscale.to(device=save_device) if exl2 else None,
        'sscale': sscale.to(device=save_device) if exl2 else None,
        'perm_inv': perm_inv.to(device=save_device),
        'group_sizes': sizes_out.to(device=save_device),
        'group_bit_widths': bits_out.to(device=save_device),
        'scale_bit_width': scale_bit_width.to(dtype=torch.uint8, device=save_device) if scale_bit_width is not None else None,
    }
    has_hook = hessian_hook is not None
    if _row_segments is None:
        if defer_metrics:
            return {'quant_meta': quant_meta, 'metrics': DeferredMetrics(stats[0], float(n_rows * n_columns), has_hook, dtype)}
        # the one host sync of the call, after every kernel of the result assembly has been queued behind the solver
        return {'quant_meta': quant_meta, 'metrics': _metrics_from_sums(stats[0].tolist(), float(n_rows * n_columns), has_hook, dtype)}
    # several linears solved as one stacked weight (gptq_quant_stacked): one result per row segment, the row-wise tensors
    # are slices of the stacked ones, everything else is shared
    host_stats = None if defer_metrics else stats.tolist()
    results, lo = [], 0
    for si, n in enumerate(seg_rows):
        meta = dict(quant_meta)
        for k in ('qweight', 'scale', 'qzero'):
            meta[k] = quant_meta[k][lo:lo + n]
        n_el = float(n * n_columns)
        met = (DeferredMetrics(stats[si], n_el, has_hook, dtype) if defer_metrics
               else _metrics_from_sums(host_stats[si], n_el, has_hook, dtype))
        results.append({'quant_meta': meta, 'metrics': met})
        lo += n
    return results


@torch.no_grad()
def gptq_quant_stacked(weights, hessian_hook, **kwargs) -> list[dict]:
    """gptq_quant of several linears that share one Hessian (q / k / v, up / gate) as ONE solve of the row-stacked weight:
    GPTQ rows are independent given U (ref :187-193), so the results are those of the separate calls, but the serial column
    recurrence runs once over 128-row tiles instead of once per linear over 32-row tiles.  Returns one {'quant_meta',
    'metrics'} dict per linear.  Not for EXL2 scale quantisation (scale_bit_width), whose grid is a maximum over all rows."""
    if kwargs.get('scale_bit_width') is not None:
        raise NotImplementedError('EXL2 scale quantisation couples the rows of a linear: quantise the linears one by one')
    weights = list(weights)
    if len(weights) == 1:
        return [gptq_quant(weight=weights[0], hessian_hook=hessian_hook, **kwargs)]
    stacked = torch.cat(weights, dim=0)
    return gptq_quant(weight=stacked, hessian_hook=hessian_hook, _row_segments=[w.shape[0] for w in weights], **kwargs)
