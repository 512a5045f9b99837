#!/usr/bin/env python
"""Timeline of the fused super-block solver's chain warp (gq_set_solve_debug): where do the ~20 us per block go?
python tools/solve_timeline.py [R] [C]"""
import ctypes
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch
import gptq
dev = torch.device('cuda')
R = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
C = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
X = torch.randn(2 * C, C, device=dev); H = X.t() @ X
_, U, _ = gptq.hessian_invert(H, None, 1.0, 0.01 * float(H.diagonal().mean()), want_hp=False)
W = (torch.randn(R, C, device=dev) * 0.02).half()
buf = torch.zeros(2048, dtype=torch.int64, device=dev)
for it in range(2):
    Wt = gptq.prep_weight_t(W, None)
    gptq.lib.gq_set_solve_debug(ctypes.c_void_p(buf.data_ptr()))
    gptq.gptq_solve(Wt, U, 128, 15.0)
    torch.cuda.synchronize()
    gptq.lib.gq_set_solve_debug(None)
allt = buf.cpu().tolist()
t0 = allt[0]
print(f'R={R} C={C}: clock64 stamps of CTA 0 (cycles since the chain warp entered block 0); reporters = warps 0 (chain), 1, 4, 15')
# per block: 4 boundary stamps; per sub-block: top, after chain, before bar1, after bar1, [before bar2 if sb < 7]
names = {0: 'chain', 1: 'w1', 2: 'w4', 3: 'w15'}
for b in range(2):
    base = b * (7 + 8 * 5 - 1)
    print(f'block {b}')
    print('   boundary (before bar, after bar, W loads issued, accfull seen, tmem read, after stage, go): ' + ' | '.join(
        names[r] + ' ' + ' '.join(str(allt[256 * r + base + k] - t0) for k in range(7)) for r in range(4)))
    for sb in range(8):
        o = base + 7 + sb * 5
        n = 5 if sb < 7 else 4
        print(f'   sb {sb}: (top, chain done, before bar1, after bar1, before bar2) ' + ' | '.join(
            names[r] + ' ' + ' '.join(str(allt[256 * r + o + k] - t0) for k in range(n)) for r in range(4)))

print('MMA issuer (warp 19): per sub-block (errors in TMEM seen, U tiles split, MMAs issued + committed)')
for u in range(16):
    print(f'   block {u // 8} sb {u % 8}: ' + ' '.join(str(allt[1024 + 3 * u + k] - t0) for k in range(3)))
