"""Multi-GPU sharding of the GPTQ hot path on one NVSwitch box (absent in the reference,
which is single-process single-GPU: main.py:25, parse_args.py:53-55).

One process per GPU (`torch.distributed`, backend nccl on the box, gloo in the CPU tests).  The path shards in the
two ways the algebra allows:

  1. H = sum over calibration sequences of X^T X  ->  rank p accumulates sequences [p*n/P, (p+1)*n/P) with K1,
     then ONE all_reduce(sum, fp32) of H's packed upper triangle per Hessian group.  `n_samples` becomes the global sequence count, so the
     2/n scale of invert() (gptq_py.py:41) is unchanged.
  2. Given U = chol(H^-1), the rows of W are independent (gptq_py.py:187-193, quant.py:114-118 are per row) ->
     rank p solves rows [p*R/P, (p+1)*R/P) of every linear of the group; codes / scales / zeros are all-gathered.
     The linears of one Hessian group are independent too ("modules" mode, the north_star wording): gptq_quant_group_sharded
     spreads a group's linears over the ranks first and shards rows only inside a linear's own ranks (plan_group),
     because the column loop of a linear is a serial chain that does not get shorter with fewer rows.

The factorisation (K2) is replicated on every rank: no broadcast of U (67-765 MB), lowest wall time.
The compute objects (`HessianHook`, `gptq_quant`) are parameters so that the CPU tests can drive the same host
logic with the oracle as the compute backend; the defaults are this build's CUDA implementations.
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence

import torch
import torch.distributed as dist


def world_info(group=None) -> tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(n: int, world: int, rank: int, multiple: int = 1) -> tuple[int, int]:
    """Contiguous shard [lo, hi) of n items for `rank`; shard boundaries are multiples of `multiple`
    (the solver needs row counts that are multiples of 4, gptq_cpp_kernel.cu:130-131).  Ragged n is allowed:
    the first ranks get one more unit, trailing ranks may be empty."""
    if n % multiple != 0:
        raise ValueError(f'{n} is not a multiple of {multiple}')
    units = n // multiple
    base, extra = divmod(units, world)
    lo = rank * base + min(rank, extra)
    hi = lo + base + (1 if rank < extra else 0)
    return lo * multiple, hi * multiple


def allreduce_hessian(hook, n_samples_global: Optional[int] = None, group=None) -> None:
    """Sum the partial Hessians of all ranks in place and make n_samples global.
    When n_samples_global is None it is all-reduced too (ragged shards)."""
    rank, world = world_info(group)
    if world == 1:
        if n_samples_global is not None:
            hook.n_samples = n_samples_global
        return
    H = hook.hessian
    if H.is_cuda and getattr(hook, 'upper_only', True) and H.shape[0] % 4 == 0:
        # K1 keeps only the upper triangle (HessianHook.add_batch, mode 0) and invert() reads nothing else: reduce the
        # upper 128 x 128 tiles as ONE contiguous message of about half the bytes (242 MB instead of 485 MB at D = 11008)
        import gptq as _ops
        packed = _ops.pack_upper(H)
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
        _ops.unpack_upper(packed, H)
    else:
        dist.all_reduce(H, op=dist.ReduceOp.SUM, group=group)
    if n_samples_global is None:
        t = torch.tensor([hook.n_samples], dtype=torch.int64, device=hook.hessian.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        n_samples_global = int(t.item())
    hook.n_samples = n_samples_global


def _all_gather_rows(local: torch.Tensor, rows_total: int, world: int, multiple: int, group=None) -> torch.Tensor:
    """Concatenate row shards produced with shard_range(rows_total, world, r, multiple)."""
    sizes = [shard_range(rows_total, world, r, multiple) for r in range(world)]
    if all(hi - lo == sizes[0][1] - sizes[0][0] for lo, hi in sizes):
        out = torch.empty((rows_total,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    # ragged: pad every shard to the largest one (uneven all_gather is not portable across backends)
    mx = max(hi - lo for lo, hi in sizes)
    padded = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * mx:r * mx + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], dim=0)


def gptq_quant_row_sharded(weight: torch.Tensor, hessian_hook, quant_fn: Optional[Callable] = None, group=None,
                           gather: bool = True, **quant_kwargs) -> dict:
    """gptq_quant of one linear with its rows sharded over the ranks.

    weight: the FULL (R, C) weight on every rank (weights are replicated; only the work is sharded).
    Returns the same {'quant_meta', 'metrics'} dict as gptq_quant; with gather=True quant_meta holds all R rows on
    every rank, metrics are the global means.  `quant_fn` defaults to this build's CUDA gptq_quant."""
    if quant_fn is None:
        from gptq_py import gptq_quant as quant_fn
    rank, world = world_info(group)
    R = weight.shape[0]
    if world == 1:
        return quant_fn(weight=weight, hessian_hook=hessian_hook, **quant_kwargs)
    if quant_kwargs.get('scale_bit_width') is not None:
        # sscale is a maximum over ALL rows of a group (quant.py:110) and reorders the stored groups (gptq_py.py:210):
        # not a row-local quantity.  Symmetric grids, the MSE search and per-group bit widths are row-local and shard.
        raise NotImplementedError('EXL2 scale quantisation does not shard by rows; quantise this linear on one rank')
    lo, hi = shard_range(R, world, rank, 4)
    dev = weight.device
    kw = dict(quant_kwargs)
    kw['save_device'] = dev                                   # gather on the compute device, move afterwards
    save_device = quant_kwargs.get('save_device', torch.device('cpu'))
    if hi > lo:
        res = quant_fn(weight=weight[lo:hi], hessian_hook=hessian_hook, **kw)
        meta, met = res['quant_meta'], res['metrics']
    else:
        meta, met = None, None
    # metrics are means over the shard's elements (gptq_py.py:196-199, 232-237): combine weighted by row count
    keys = ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')
    m = torch.tensor([(met[k] if met is not None else 0.0) * (hi - lo) for k in keys], dtype=torch.float64, device=dev)
    dist.all_reduce(m, group=group)
    metrics = {k: float(v) / R for k, v in zip(keys, m.tolist())}
    if not gather:
        return {'quant_meta': meta, 'metrics': metrics, 'rows': (lo, hi)}
    C = weight.shape[1]
    # every rank derives the same G / dtypes even when its own shard is empty
    gs = quant_kwargs['group_sizes']
    G = len(gs)
    if meta is None:
        meta = {'qweight': torch.empty(0, C, dtype=torch.uint8, device=dev),
                'scale': torch.empty(0, G, dtype=weight.dtype, device=dev),
                'qzero': torch.empty(0, G, dtype=torch.uint8, device=dev)}
    full = {}
    for k in ('qweight', 'scale', 'qzero'):
        full[k] = _all_gather_rows(meta[k].to(dev), R, world, 4, group).to(save_device)
    perm = getattr(hessian_hook, 'perm', None)
    if perm is not None:
        perm_inv = perm.to(torch.int16).argsort().to(torch.int16)
    else:
        perm_inv = torch.arange(C, dtype=torch.int16, device=dev)
    full.update({'qscale': None, 'sscale': None, 'perm_inv': perm_inv.to(save_device),
                 'group_sizes': gs.to(dtype=torch.int16, device=save_device),
                 'group_bit_widths': quant_kwargs['group_bit_widths'].to(dtype=torch.uint8, device=save_device),
                 'scale_bit_width': None})
    return {'quant_meta': full, 'metrics': metrics}


def plan_group(rows: Sequence[int], world: int, multiple: int = 4) -> list[tuple[int, int, int, int]]:
    """Work plan for the L linears of ONE Hessian group (they share U and are independent of each other) on `world`
    ranks: a list of (linear index, row lo, row hi, rank).  The column loop of a linear is a serial chain whose length
    does not shrink with the rows (C / 128 dependent block solves), so linears are spread over ranks FIRST and rows
    only inside a linear's own ranks: with world >= L every linear gets a contiguous set of ranks (sizes as even as
    possible, extra ranks to the linears with more rows) and its rows are split over them with shard_range; with
    world < L whole linears are dealt to the least-loaded rank, longest first.  Deterministic, identical on every rank."""
    L = len(rows)
    plan: list[tuple[int, int, int, int]] = []
    if L == 0:
        return plan
    if world >= L:
        base, extra = divmod(world, L)
        order = sorted(range(L), key=lambda i: (-rows[i], i))
        n_ranks = [base] * L
        for i in order[:extra]:
            n_ranks[i] += 1
        r0 = 0
        for i in range(L):
            for j in range(n_ranks[i]):
                lo, hi = shard_range(rows[i], n_ranks[i], j, multiple)
                if hi > lo:
                    plan.append((i, lo, hi, r0 + j))
            r0 += n_ranks[i]
    else:
        load = [0] * world
        for i in sorted(range(L), key=lambda i: (-rows[i], i)):
            r = min(range(world), key=lambda q: (load[q], q))
            plan.append((i, 0, rows[i], r))
            load[r] += rows[i]
        plan.sort()
    return plan


class _Unchecked:
    """View of a HessianHook whose deferred check is still pending, for the work that is queued before settle()."""

    def __init__(self, hook):
        self._hook = hook

    def __getattr__(self, name):
        if name == '_pending':
            return None
        return getattr(self._hook, name)


def _solve_cost(rows: int, n_sms: int = 148) -> float:
    """Relative time of the fused solver for one weight of `rows` rows (measured on B200, profiles/r02_v24_quant_time.jsonl:
    32-row CTAs 1.0, 64-row 1.5, 96-row 1.9, 128-row 2.1 per wave; the library picks the smallest tile that fits one wave)."""
    for tile, cost in ((32, 1.0), (64, 1.5), (96, 1.9), (128, 2.1)):
        if -(-rows // tile) <= n_sms:
            return cost
    return 2.1 * -(-(-(-rows // 128)) // n_sms)


def _stacking_pays(rows: Sequence[int]) -> bool:
    """Solve the linears of a Hessian group as one row-stacked weight?  (7B q/k/v: 3 x 1.0 -> 1.95; up/gate: 2 x 1.95 -> 3.9, no.)"""
    return len(rows) > 1 and len(rows) <= 8 and all(r % 4 == 0 for r in rows) and _solve_cost(sum(rows)) < 0.95 * sum(_solve_cost(r) for r in rows)


def gptq_quant_group_sharded(weights: Sequence[torch.Tensor], hessian_hook, quant_fn: Optional[Callable] = None, group=None,
                             shard_of: Optional[Callable[[int, int, int], torch.Tensor]] = None,
                             **quant_kwargs) -> list[dict]:
    """gptq_quant of all linears that share one Hessian (q/k/v, up/gate, or a single o / down), spread over the ranks
    by plan_group.  weights: the FULL (R_i, C) weights, replicated on every rank.  Returns one {'quant_meta',
    'metrics'} dict per linear, complete on every rank: the shards travel as ONE all_gather of a per-rank byte blob
    (codes, scales and zero points of everything the rank solved, padded to the largest blob) and one all_reduce of
    the metric sums.
    shard_of(i, lo, hi): optional source of rows [lo, hi) of linear i for callers that stage only the rows a rank solves
    (bench.py's host-fed step); `weights` may then be meta tensors (shape and dtype only)."""
    own_quant = quant_fn is None
    if own_quant:
        from gptq_py import gptq_quant as quant_fn
        # the library's own quantiser: no host synchronisation per linear, the metrics are read when the caller looks
        quant_kwargs = dict(quant_kwargs, defer_metrics=True)
    rank, world = world_info(group)
    L = len(weights)
    if shard_of is None:
        def shard_of(i, lo, hi):
            return weights[i][lo:hi]
    if own_quant and getattr(hessian_hook, '_pending', None) is not None:
        # HessianHook.invert(defer_check=True): the group is queued behind the unchecked factorisation, the check (one host
        # synchronisation) comes after; if the first try was bad (rare: a damping retry), everything is redone with the final U
        first = gptq_quant_group_sharded(weights, _Unchecked(hessian_hook), None, group, shard_of, **{k: v for k, v in quant_kwargs.items() if k != 'defer_metrics'})
        if hessian_hook.settle():
            return first
        return gptq_quant_group_sharded(weights, hessian_hook, None, group, shard_of, **{k: v for k, v in quant_kwargs.items() if k != 'defer_metrics'})
    if world == 1:
        if own_quant and _stacking_pays([w.shape[0] for w in weights]) and quant_kwargs.get('scale_bit_width') is None \
                and len({(w.shape[1], w.dtype) for w in weights}) == 1:
            # the linears share U and GPTQ rows are independent: one solve of the row-stacked weight (128-row tiles, one wave)
            # instead of one serial column recurrence per linear
            from gptq_py import gptq_quant_stacked
            return gptq_quant_stacked([shard_of(i, 0, w.shape[0]) for i, w in enumerate(weights)], hessian_hook, **quant_kwargs)
        return [quant_fn(weight=shard_of(i, 0, w.shape[0]), hessian_hook=hessian_hook, **quant_kwargs)
                for i, w in enumerate(weights)]
    if quant_kwargs.get('scale_bit_width') is not None:
        raise NotImplementedError('EXL2 scale quantisation does not shard by rows; quantise this linear on one rank')
    dev = hessian_hook.hessian_inv.device
    dtype = weights[0].dtype
    C = weights[0].shape[1]
    assert all(w.shape[1] == C and w.dtype == dtype for w in weights)
    gs = quant_kwargs['group_sizes']
    G = len(gs)
    save_device = quant_kwargs.get('save_device', torch.device('cpu'))
    kw = dict(quant_kwargs)
    kw['save_device'] = dev
    plan = plan_group([w.shape[0] for w in weights], world, 4)
    esz = torch.empty(0, dtype=dtype).element_size()
    fields = (('qweight', C, torch.uint8), ('scale', G * esz, dtype), ('qzero', G, torch.uint8))   # bytes per row

    def padded(nbytes: int) -> int:
        return (nbytes + 15) & ~15                              # every field starts 16-byte aligned (typed views)

    def shard_bytes(n: int) -> int:
        return sum(padded(n * width) for _, width, _ in fields)

    blob_bytes = [0] * world
    for i, lo, hi, r in plan:
        blob_bytes[r] += shard_bytes(hi - lo)
    mx = max(max(blob_bytes), 1)
    blob = torch.zeros(mx, dtype=torch.uint8, device=dev)
    keys = ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')
    msum = torch.zeros(L, len(keys), dtype=torch.float64, device=dev)
    off = 0
    mine = [(i, lo, hi) for i, lo, hi, r in plan if r == rank]
    if own_quant and _stacking_pays([hi - lo for _, lo, hi in mine]):
        # this rank's shards of the group (rows of different linears, same U) as one row-stacked solve
        from gptq_py import gptq_quant_stacked
        mine_res = gptq_quant_stacked([shard_of(i, lo, hi) for i, lo, hi in mine], hessian_hook, **kw)
    else:
        mine_res = [quant_fn(weight=shard_of(i, lo, hi), hessian_hook=hessian_hook, **kw) for i, lo, hi in mine]
    for (i, lo, hi), res in zip(mine, mine_res):
        meta, met = res['quant_meta'], res['metrics']
        for k, width, _ in fields:
            b = meta[k].to(dev).contiguous().view(torch.uint8).reshape(-1)
            assert b.numel() == (hi - lo) * width
            blob[off:off + b.numel()] = b
            off += padded(b.numel())
        if hasattr(met, 'device_values'):
            msum[i] += met.device_values() * float(hi - lo)            # stays on the device (keys order == DeferredMetrics.KEYS)
        else:
            msum[i] += torch.tensor([met[k] * (hi - lo) for k in keys], dtype=torch.float64, device=dev)
    out = torch.empty(world * mx, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, blob, group=group)
    dist.all_reduce(msum, group=group)
    # reassemble: every rank walks the same plan
    pieces: list[dict[str, list]] = [{'qweight': [], 'scale': [], 'qzero': []} for _ in range(L)]
    offs = [0] * world
    for i, lo, hi, r in plan:
        n = hi - lo
        base = r * mx + offs[r]
        for k, width, dt in fields:
            chunk = out[base:base + n * width]
            pieces[i][k].append(chunk.view(dt).reshape(n, -1) if dt != torch.uint8 else chunk.reshape(n, -1))
            base += padded(n * width)
        offs[r] += shard_bytes(n)
    perm = getattr(hessian_hook, 'perm', None)
    if perm is not None:
        perm_inv = perm.to(torch.int16).argsort().to(torch.int16)
    else:
        perm_inv = torch.arange(C, dtype=torch.int16, device=dev)
    msum = msum.tolist()
    results = []
    for i in range(L):
        R = weights[i].shape[0]
        meta = {k: torch.cat(v, dim=0).to(save_device) for k, v in pieces[i].items()}
        meta.update({'qscale': None, 'sscale': None, 'perm_inv': perm_inv.to(save_device),
                     'group_sizes': gs.to(dtype=torch.int16, device=save_device),
                     'group_bit_widths': quant_kwargs['group_bit_widths'].to(dtype=torch.uint8, device=save_device),
                     'scale_bit_width': None})
        results.append({'quant_meta': meta, 'metrics': {k: v / R for k, v in zip(keys, msum[i])}})
    return results
