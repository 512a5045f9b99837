#!/usr/bin/env python
"""K2 in isolation: HessianHook.invert on a synthetic Hessian of size D (16 x 2048 tokens, log-uniform column scales), median of
5 CUDA-event timings with programmatic dependent launch on and off, and the distance of U from the float64 factor.
    python tools/invert_time.py 4096 11008"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402


def main():
    import gptq
    from gptq_py import HessianHook
    dev = torch.device('cuda:0')
    for D in [int(a) for a in sys.argv[1:]] or [4096, 11008]:
        g = torch.Generator(device=dev).manual_seed(D)
        cs = torch.exp(torch.rand(D, device=dev, generator=g) * 3.0 - 1.5)
        X = (torch.randn(16 * 2048, D, device=dev, generator=g) * cs).half()
        hook = HessianHook()
        hook.add_batch(X.view(16, 2048, D))
        del X
        H0 = hook.hessian.clone()
        out = {'D': D}
        U64 = None
        for pdl in (1, 0, 1):
            gptq.set_pdl(pdl)
            ts = []
            for rep in range(6):
                h = HessianHook()
                h.hessian, h.n_samples = H0.clone(), 16
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                h.invert(1e-2, True)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            if U64 is None:
                Hp = h.hessian.double()
                L = torch.linalg.cholesky(Hp)
                U64 = torch.linalg.cholesky(torch.cholesky_inverse(L), upper=True)
                del Hp, L
            out[f'pdl={pdl}' + ('' if f'pdl={pdl}' not in out else ' (again)')] = {
                'invert_ms_median': round(sorted(ts[1:])[2], 3), 'min': round(min(ts[1:]), 3),
                'U_rel_fro_vs_fp64': float((h.hessian_inv.double() - U64).norm() / U64.norm()), 'tries': h.invert_tries}
        gptq.set_pdl(1)
        print(json.dumps(out))
        del U64


if __name__ == '__main__':
    main()
