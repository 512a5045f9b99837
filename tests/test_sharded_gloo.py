"""N > 1 host logic (gptq-v2_b200/sharded.py) on CPU: world_size 2, gloo, with the ORACLE as the compute backend
(injected through sharded.py's quant_fn / hook parameters; the product defaults are the CUDA ops).

Checked: sequence-sharded partial Hessians + all_reduce == the single-process Hessian; row-sharded solve + all_gather
== the single-process result bit for bit (rows are independent given U); ragged shards; metrics recombination."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


class OracleHook:
    """HessianHook look-alike (fields of gptq_py.py:10-16) computing with the CPU oracle."""

    def __init__(self):
        self.hessian, self.n_samples, self.hessian_inv, self.perm, self.perm_inv = None, 0, None, None, None

    def add_batch(self, inp):
        from oracle import gptq_oracle as O
        self.hessian, self.n_samples = O.hessian_accumulate(self.hessian, self.n_samples, inp)

    def invert(self, damp_ratio=1e-2, act_order=True):
        from oracle import gptq_oracle as O
        inv = O.hessian_invert(self.hessian, self.n_samples, damp_ratio, act_order)
        self.hessian, self.hessian_inv, self.perm, self.perm_inv = inv['H'], inv['U'], inv['perm'], inv['perm_inv']
        return self.hessian_inv


def oracle_quant(weight, hessian_hook, group_sizes, group_bit_widths, save_device=torch.device('cpu'), **_):
    from oracle import gptq_oracle as O
    gs, bits = int(group_sizes[0]), int(group_bit_widths[0])
    sol = O.gptq_solve_pure(weight, hessian_hook.hessian_inv, hessian_hook.perm, gs, bits)
    return O.assemble_result(weight, sol, hessian_hook.hessian_inv, hessian_hook.perm, gs, bits)


def _worker(rank, world, port, R, C, n_seq, ret):
    for p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200'), os.path.join(ROOT, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.set_num_threads(2)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import sharded
        from oracle import gptq_oracle as O
        W, X = O.synth_linear(R, C, n_seq, 64, seed=5)
        lo, hi = sharded.shard_range(n_seq, world, rank)
        hook = OracleHook()
        for i in range(lo, hi):
            hook.add_batch(X[i:i + 1])
        if hi == lo:                                   # a rank without sequences still takes part in the collective
            hook.hessian = torch.zeros(C, C)
        sharded.allreduce_hessian(hook, None)
        assert hook.n_samples == n_seq
        H_sum = hook.hessian.clone()
        hook.invert(1e-2, True)
        gs = torch.full([C // 128], 128, dtype=torch.int32)
        bw = torch.full([C // 128], 4, dtype=torch.int32)
        res = sharded.gptq_quant_row_sharded(W, hook, quant_fn=oracle_quant, group_sizes=gs, group_bit_widths=bw)
        # EXL2 scale quantisation needs a maximum over all rows of a group: the row-sharded path must refuse it
        try:
            sharded.gptq_quant_row_sharded(W, hook, quant_fn=oracle_quant, group_sizes=gs, group_bit_widths=bw,
                                           scale_bit_width=torch.tensor(4))
            raise AssertionError('EXL2 was accepted by the row-sharded path')
        except NotImplementedError:
            pass
        if rank == 0:
            torch.save({'H': H_sum, 'U': hook.hessian_inv, 'perm': hook.perm, 'meta': res['quant_meta'],
                        'metrics': res['metrics']}, ret)
        dist.barrier()
    except Exception:
        import traceback
        with open(ret + f'.err{rank}', 'w') as f:
            f.write(traceback.format_exc())
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('R,C,n_seq', [(64, 256, 6), (36, 128, 3)])
def test_sequence_and_row_sharding_world2(R, C, n_seq, tmp_path):
    from oracle import gptq_oracle as O
    ctx = mp.get_context('spawn')
    ret = os.path.join(str(tmp_path), 'rank0.pt')               # results come back through a file
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, R, C, n_seq, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        if p.is_alive():
            p.kill()
    errs = [open(ret + f'.err{r}').read() for r in range(2) if os.path.exists(ret + f'.err{r}')]
    assert [p.exitcode for p in procs] == [0, 0], ([p.exitcode for p in procs], errs)
    got = torch.load(ret)
    # single-process oracle on the same inputs
    W, X = O.synth_linear(R, C, n_seq, 64, seed=5)
    H, n = None, 0
    for i in range(n_seq):
        H, n = O.hessian_accumulate(H, n, X[i:i + 1])
    assert O.rel_fro(got['H'], H) <= 1e-6                       # same terms, different summation order
    # given the all-reduced H, the row-sharded solve must equal the unsharded one bit for bit
    sol = O.gptq_solve_pure(W, got['U'], got['perm'], 128, 4)
    want = O.assemble_result(W, sol, got['U'], got['perm'], 128, 4)
    for k in ('qweight', 'scale', 'qzero', 'perm_inv', 'group_sizes', 'group_bit_widths'):
        assert torch.equal(got['meta'][k], want['quant_meta'][k]), k
    for k, v in want['metrics'].items():
        # l2.0_norm is a mean taken in the weight dtype (fp16, gptq_py.py:237): shards round differently
        assert got['metrics'][k] == pytest.approx(v, rel=2e-3 if k == 'l2.0_norm' else 1e-5), k
    # and the end result agrees with the fully single-process run (H differs only by summation order)
    full = O.run_linear(W, X, 128, 4, 1e-2, True, mode='pure', batch=1)
    assert O.code_match(got['meta']['qweight'], full['result']['quant_meta']['qweight']) >= 0.999


def _group_worker(rank, world, port, rows, C, ret):
    for p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200'), os.path.join(ROOT, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.set_num_threads(2)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import sharded
        from oracle import gptq_oracle as O
        _, X = O.synth_linear(8, C, 4, 64, seed=9)
        hook = OracleHook()
        hook.add_batch(X)
        hook.invert(1e-2, True)
        Ws = [O.synth_linear(R, C, 1, 8, seed=20 + i)[0] for i, R in enumerate(rows)]
        gs = torch.full([C // 128], 128, dtype=torch.int32)
        bw = torch.full([C // 128], 4, dtype=torch.int32)
        res = sharded.gptq_quant_group_sharded(Ws, hook, quant_fn=oracle_quant, group_sizes=gs, group_bit_widths=bw)
        assert len(res) == len(rows)
        for W, r in zip(Ws, res):                              # every rank holds every linear, bit for bit
            want = oracle_quant(W, hook, gs, bw)
            for k in ('qweight', 'scale', 'qzero', 'perm_inv', 'group_sizes', 'group_bit_widths'):
                assert torch.equal(r['quant_meta'][k], want['quant_meta'][k]), (rank, k)
            for k, v in want['metrics'].items():
                assert r['metrics'][k] == pytest.approx(v, rel=2e-3 if k == 'l2.0_norm' else 1e-5), (rank, k)
        if rank == 0:
            torch.save({'ok': True}, ret)
        dist.barrier()
    except Exception:
        import traceback
        with open(ret + f'.err{rank}', 'w') as f:
            f.write(traceback.format_exc())
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('rows', [(64, 64, 64), (36,), (40, 24)])
def test_group_sharding_world2(rows, tmp_path):
    """Linears of one Hessian group spread over ranks (whole linears when world < L, rows inside a linear's own ranks
    otherwise), one blob all_gather: every rank ends with every linear's full quant_meta, identical to the
    single-process result."""
    ctx = mp.get_context('spawn')
    ret = os.path.join(str(tmp_path), 'rank0.pt')
    port = _free_port()
    procs = [ctx.Process(target=_group_worker, args=(r, 2, port, rows, 128, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        if p.is_alive():
            p.kill()
    errs = [open(ret + f'.err{r}').read() for r in range(2) if os.path.exists(ret + f'.err{r}')]
    assert [p.exitcode for p in procs] == [0, 0], ([p.exitcode for p in procs], errs)
    assert torch.load(ret)['ok']


def test_plan_group_properties():
    import sharded
    for rows, world in (([4096] * 3, 8), ([4096] * 3, 2), ([11008] * 2, 8), ([4096], 8), ([4096] * 3, 4), ([4096, 1024, 1024], 8),
                        ([12] * 3, 8), ([5120] * 3, 1), ([8], 4)):
        plan = sharded.plan_group(rows, world, 4)
        cover = [[] for _ in rows]
        for i, lo, hi, r in plan:
            assert 0 <= r < world and 0 <= lo < hi <= rows[i] and lo % 4 == 0 and hi % 4 == 0
            cover[i].append((lo, hi))
        for i, R in enumerate(rows):                           # every row of every linear exactly once
            segs = sorted(cover[i])
            assert segs[0][0] == 0 and segs[-1][1] == R
            assert all(a[1] == b[0] for a, b in zip(segs, segs[1:]))
        if world >= len(rows):                                 # a rank never works on two linears
            owners = {}
            for i, lo, hi, r in plan:
                assert owners.setdefault(r, i) == i


def test_shard_range_properties():
    import sharded
    for n, world, mult in ((128, 8, 1), (4096, 8, 4), (11008, 8, 4), (36, 2, 4), (12, 8, 4), (0, 4, 1)):
        parts = [sharded.shard_range(n, world, r, mult) for r in range(world)]
        assert parts[0][0] == 0 and parts[-1][1] == n
        for (a, b), (c, d) in zip(parts, parts[1:]):
            assert b == c and a <= b and (b - a) % mult == 0
        sizes = [b - a for a, b in parts]
        assert max(sizes) - min(sizes) <= mult
    with pytest.raises(ValueError):
        sharded.shard_range(10, 2, 0, 4)
