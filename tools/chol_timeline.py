#!/usr/bin/env python
"""clock64 timeline of the first chol_panel_kernel launch of a factorisation (CTA 0, warps 0 and 1): python tools/chol_timeline.py [D]"""
import ctypes
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch
import gptq
dev = torch.device('cuda')
D = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
X = torch.randn(2 * D, D, device=dev); H = X.t() @ X
buf = torch.zeros(256, dtype=torch.int64, device=dev)
for it in range(2):
    gptq.lib.gq_set_solve_debug(ctypes.c_void_p(buf.data_ptr()))
    gptq.hessian_invert(H, None, 1.0, 0.01 * float(H.diagonal().mean()), want_hp=False)
    torch.cuda.synchronize()
    gptq.lib.gq_set_solve_debug(None)
t = buf.cpu().tolist()
t0 = t[0]
print(f'D={D}: chol_panel_kernel<64>, step 0, CTA 0; cycles since kernel entry; columns: warp 0 | warp 1')
labels = ['entry', 'staged', 'factor(0) done', 'sync']
for s in range(8):
    labels += [f's{s} B done']
    if s < 7:
        labels += [f's{s} look-ahead chunk done', f's{s} factor done (warp 0)', f's{s} updates done', f's{s} sync']
labels += ['written back']
for i, name in enumerate(labels):
    print(f'  {name:28s} {t[i] - t0:8d} | {t[64 + i] - t0:8d}')
