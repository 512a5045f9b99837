#!/usr/bin/env python
"""Small fixed workloads for ncu: python tools/prof_targets.py {tf32|syrk|invert|solve|w4a16}"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch
import gptq
dev = torch.device('cuda')
which = sys.argv[1]
if os.environ.get('GQ_TF32_MODE'):
    gptq.set_tf32x3(int(os.environ['GQ_TF32_MODE']))
if which == 'tf32':
    # default: one K4 near update; `tf32 9984 4096 1024` = the far update behind a super-block of 8 blocks (down_proj)
    M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (3968, 4096, 128)
    A = torch.randn(K, M, device=dev); B = torch.randn(K, N, device=dev); C = torch.zeros(M, N, device=dev)
    for _ in range(3):
        gptq.gemm_tf32x3(A, B, C, M, N, K, M, N, N, -1.0, False)
elif which == 'syrk':
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    X = torch.randn(32768, D, device=dev).half(); H = torch.zeros(D, D, device=dev)
    for _ in range(3):
        gptq.hessian_syrk(H, X, mode=0, variant=2)
elif which == 'invert':
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    X = torch.randn(2 * D, D, device=dev); H = X.t() @ X
    for _ in range(2):
        gptq.hessian_invert(H, None, 1.0, 0.01 * float(H.diagonal().mean()))
elif which == 'solve':
    R, C = 4096, 4096
    X = torch.randn(2 * C, C, device=dev); H = X.t() @ X
    _, U, _ = gptq.hessian_invert(H, None, 1.0, 0.01 * float(H.diagonal().mean()), want_hp=False)
    W = (torch.randn(R, C, device=dev) * 0.02).half()
    for _ in range(2):
        gptq.gptq_solve(gptq.prep_weight_t(W, None), U, 128, 15.0)
elif which == 'w4a16':
    N, K = 28672, 8192
    q = torch.randint(-2 ** 31, 2 ** 31 - 1, (N, K // 8), device=dev, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, device=dev) * 0.01).half(); qz = torch.randint(0, 16, (K // 128, N), device=dev, dtype=torch.uint8)
    x = torch.randn(1, K, device=dev).half()
    for _ in range(3):
        gptq.w4a16_gemm(x, q, sc, qz, None, 128)
torch.cuda.synchronize()
print('ok')
