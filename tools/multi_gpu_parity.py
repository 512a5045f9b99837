#!/usr/bin/env python
"""N-rank result == 1-rank result: run under torchrun on N GPUs of one box.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/multi_gpu_parity.py [--small]

One Llama-2-7B decoder layer's hot path (bench.py's step: 4 Hessian groups, 7 linears, 128 x 2048 tokens) is run twice
by the SAME processes on the SAME synthetic data:
  sharded : sequences split over the ranks -> K1 per rank -> NCCL all-reduce of the packed upper triangle of H ->
            K2 -> the group's linears / rows spread over the ranks (sharded.plan_group) -> one all-gather per group
  single  : rank 0 alone, all sequences, plain HessianHook / gptq_quant calls
and rank 0 asserts, per linear:
  (1) with the all-reduced H injected into the single-rank path (same H, same U), qweight / scale / qzero of the sharded
      path are IDENTICAL to the single-rank result: row sharding, the linear-to-rank plan and the blob all-gather change
      nothing (rows are independent given U, /root/reference/gptq_py.py:187-193);
  (2) H summed over ranks vs H accumulated by one rank: relative Frobenius distance <= 1e-6 (fp32 summation order of
      the partial Hessians), codes of the fully independent single-rank pipeline >= 99.9 % equal, reconstruction-error
      metric within 1 %.
Prints one JSON line (kept under profiles/)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--small', action='store_true', help='16 sequences of 512 tokens (quick check)')
    args = ap.parse_args()
    import bench as B
    import sharded
    from gptq_py import HessianHook, gptq_quant
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    dev = torch.device('cuda', local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    n_seq, seqlen = (16, 512) if args.small else (B.N_SEQ, B.SEQLEN)
    call = min(B.CALL_SEQ, n_seq // world)
    g = torch.Generator(device=dev).manual_seed(1234)                       # identical on every rank
    W = {n: (torch.randn(R, C, device=dev, generator=g) * 0.02).half() for n, R, C, _ in B.LINEARS}
    cs = {D: torch.exp(torch.rand(D, device=dev, generator=g) * 3.0 - 1.5) for D in sorted(set(B.GROUP_D))}
    report = {'world': world, 'n_seq': n_seq, 'seqlen': seqlen, 'linears': {}}
    ok = True
    for gi, D in enumerate(B.GROUP_D):
        grp = [l for l in B.LINEARS if l[3] == gi]
        gx = torch.Generator(device=dev).manual_seed(77 + gi)               # identical on every rank: all sequences
        X = torch.empty(n_seq, seqlen, D, device=dev, dtype=torch.float16)
        for i in range(n_seq):
            X[i] = (torch.randn(seqlen, D, device=dev, generator=gx) * cs[D]).half()
        C = grp[0][2]
        gs = torch.full([C // B.GROUP], B.GROUP, dtype=torch.int32)
        bw = torch.full([C // B.GROUP], B.BITS, dtype=torch.int32)
        kw = dict(group_sizes=gs, group_bit_widths=bw, scale_bit_width=None, gptq_use_kernel=False, gptq_block_sizes=gs,
                  quant_symmetric=False, quant_mse=False, save_device=dev)
        # ---- sharded
        lo, hi = sharded.shard_range(n_seq, world, rank)
        hook = HessianHook()
        for i in range(lo, hi, call):
            hook.add_batch(X[i:min(i + call, hi)])
        sharded.allreduce_hessian(hook, n_seq)
        H_sum = hook.hessian.clone()
        hook.invert(1e-2, True)
        res_sh = sharded.gptq_quant_group_sharded([W[l[0]] for l in grp], hook, **kw)
        if rank == 0:
            # ---- single rank, same H (injected): must be identical
            h1 = HessianHook()
            h1.hessian, h1.n_samples = H_sum.clone(), n_seq
            h1.invert(1e-2, True)
            assert torch.equal(h1.hessian_inv, hook.hessian_inv) and torch.equal(h1.perm, hook.perm)
            # ---- single rank, own accumulation over all sequences
            h2 = HessianHook()
            for i in range(0, n_seq, call):
                h2.add_batch(X[i:i + call])
            H_one = torch.triu(h2.hessian)
            relH = float((torch.triu(H_sum) - H_one).double().norm() / H_one.double().norm())
            h2.invert(1e-2, True)
            same_perm = bool(torch.equal(h2.perm, hook.perm))
            for (name, R, C, _), rs in zip(grp, res_sh):
                r1 = gptq_quant(weight=W[name], hessian_hook=h1, **kw)
                r2 = gptq_quant(weight=W[name], hessian_hook=h2, **kw)
                ident = all(torch.equal(rs['quant_meta'][k], r1['quant_meta'][k]) for k in ('qweight', 'scale', 'qzero', 'perm_inv'))
                match = float((rs['quant_meta']['qweight'] == r2['quant_meta']['qweight']).float().mean()) if same_perm else None
                met = {k: (rs['metrics'][k], r1['metrics'][k], r2['metrics'][k]) for k in ('gptq_error', 'l2.0_error')}
                good = ident and relH <= 1e-6 and (match is None or match >= 0.999) and \
                    abs(met['gptq_error'][0] - met['gptq_error'][2]) <= 0.01 * met['gptq_error'][2] and \
                    abs(met['gptq_error'][0] - met['gptq_error'][1]) <= 1e-6 * met['gptq_error'][1]
                ok = ok and good
                report['linears'][name] = {'identical_given_same_H': ident, 'H_rel_fro_vs_one_rank': relH, 'same_perm': same_perm,
                                           'code_match_vs_one_rank_pipeline': match, 'gptq_error(sharded, same H, one rank)': met['gptq_error'],
                                           'ok': good}
        del X, hook
        if world > 1:
            dist.barrier()
    if rank == 0:
        report['ok'] = ok
        print(json.dumps(report))
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not ok:
        sys.exit(1)


if __name__ == '__main__':
    main()
