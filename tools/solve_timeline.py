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
buf = torch.zeros(512, dtype=torch.int64, device=dev)
for it in range(2):
    Wt = gptq.prep_weight_t(W, None)
    gptq.lib.gq_set_solve_debug(ctypes.c_void_p(buf.data_ptr()))
    gptq.gptq_solve(Wt, U, 128, 15.0)
    torch.cuda.synchronize()
    gptq.lib.gq_set_solve_debug(None)
t = buf.cpu().tolist()
i = 0
print(f'R={R} C={C}: clock64 deltas of CTA 0 / warp 0 (cycles)')
for b in range(4):
    tr = t[i:i + 4]; i += 4                       # before bar, after bar, after fill, go
    print(f'block {b}: wait-for-all {tr[1] - tr[0]:6d}  stage W/acc {tr[2] - tr[1]:6d}  params+Upk+bars {tr[3] - tr[2]:6d}')
    tot_chain = tot_pub = tot_next = 0
    start = tr[3]
    for sb in range(8):
        s0, s1, s2 = t[i], t[i + 1], t[i + 2]; i += 3          # top, chain done, before bar
        s3 = t[i]; i += 1                                       # after bar 1
        chain, tm, bar1 = s1 - s0, s2 - s1, s3 - s2
        nxt = (t[i] - s3) if sb < 7 else 0                      # next top (update + bar 2)
        tot_chain += chain; tot_pub += tm + bar1; tot_next += nxt
        print(f'   sb {sb}: chain {chain:5d}  tmem-st {tm:5d}  bar1 {bar1:5d}  next16+bar2 {nxt:5d}')
    print(f'   block total {t[i - 1] - tr[0]:6d} cycles: chain {tot_chain}, publish {tot_pub}, look-ahead {tot_next}')
