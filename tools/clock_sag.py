#!/usr/bin/env python
"""Why are K2 / K3 slower inside bench.py's step than alone?  Times HessianHook.invert and one gptq_quant
  (a) on an idle GPU, (b) queued right behind ~35 ms of K1 (the step's own situation: the board is at its power cap and
  the SM clock has dropped), (c) queued behind an idle gap of the same length.  CUDA events, 5 repetitions each, median.
    python tools/clock_sag.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402


def main():
    from gptq_py import HessianHook, gptq_quant
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    heat_x = torch.randn(16, 2048, 11008, device=dev, generator=g).half()
    heat = HessianHook()
    heat.add_batch(heat_x)

    def heat_up(calls=8):
        for _ in range(calls):
            heat.add_batch(heat_x)

    for D, R in ((4096, 4096), (11008, 4096)):
        cs = torch.exp(torch.rand(D, device=dev, generator=g) * 3.0 - 1.5)
        X = (torch.randn(16, 2048, D, device=dev, generator=g) * cs).half()
        hook = HessianHook()
        hook.add_batch(X)
        H0 = hook.hessian.clone()
        W = (torch.randn(R, D, device=dev, generator=g) * 0.02).half()
        gs = torch.full([D // 128], 128, dtype=torch.int32)
        bw = torch.full([D // 128], 4, dtype=torch.int32)
        out = {'D': D, 'R': R}
        for mode in ('idle', 'behind_k1', 'idle', 'behind_k1'):
            ti, tq = [], []
            for rep in range(6):
                h = HessianHook()
                h.hessian, h.n_samples = H0.clone(), 16
                torch.cuda.synchronize()
                time.sleep(0.05)
                if mode == 'behind_k1':
                    heat_up()
                a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                a.record()
                h.invert(1e-2, True, defer_check=True)
                b.record()
                gptq_quant(W, h, gs, bw, gptq_use_kernel=False, gptq_block_sizes=gs, quant_mse=False, save_device=dev)
                c.record()
                torch.cuda.synchronize()
                ti.append(a.elapsed_time(b))
                tq.append(b.elapsed_time(c))
            key = mode if mode not in out else mode + ' (again)'
            out[key] = {'invert_ms': round(sorted(ti[1:])[2], 3), 'gptq_quant_ms': round(sorted(tq[1:])[2], 3)}
        print(json.dumps(out), flush=True)


if __name__ == '__main__':
    main()
