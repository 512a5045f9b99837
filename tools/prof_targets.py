#!/usr/bin/env python
"""Small fixed workloads for ncu: python tools/prof_targets.py {tf32|syrk|invert|solve|w4a16|w4a16mma|gemmnt|layer}"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch
import gptq
dev = torch.device('cuda')
which = sys.argv[1]
if os.environ.get('W4_PATH_FORCE'):
    gptq.set_w4a16_path(int(os.environ['W4_PATH_FORCE']))
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
    M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (1, 28672, 8192)     # batch 1..8 / 9..32: the mma.sync kernels
    q = torch.randint(-2 ** 31, 2 ** 31 - 1, (N, K // 8), device=dev, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, device=dev) * 0.01).half(); qz = torch.randint(0, 16, (K // 128, N), device=dev, dtype=torch.uint8)
    x = torch.randn(M, K, device=dev).half()
    for _ in range(3):
        gptq.w4a16_gemm(x, q, sc, qz, None, 128)
elif which == 'w4a16mma':
    # tcgen05 TS-mode kernel (batch 64) and the dense large-batch route's own GEMM
    N, K, M = 8192, 8192, 64
    q = torch.randint(-2 ** 31, 2 ** 31 - 1, (N, K // 8), device=dev, dtype=torch.int32)
    sc = (torch.rand(K // 128, N, device=dev) * 0.01).half(); qz = torch.randint(0, 16, (K // 128, N), device=dev, dtype=torch.uint8)
    x = torch.randn(M, K, device=dev).half()
    for _ in range(3):
        gptq.w4a16_gemm(x, q, sc, qz, None, 128)
elif which == 'gemmnt':
    M, N, K = 4096, 8192, 8192
    x = torch.randn(M, K, device=dev).half(); w = torch.randn(N, K, device=dev).half()
    for _ in range(3):
        gptq.gemm_f16_nt(x, w)
elif which == 'layer':
    # every kernel of the quantisation of one linear of shape (R, C) with act-order: gather / factor / solve / pack
    R, C = (int(v) for v in sys.argv[2:4]) if len(sys.argv) >= 4 else (11008, 4096)
    from gptq_py import HessianHook, gptq_quant
    g = torch.Generator(device=dev).manual_seed(1)
    cs = torch.exp(torch.rand(C, device=dev, generator=g) * 3.0 - 1.5)
    X = (torch.randn(max(4, C // 768), 2048, C, device=dev, generator=g) * cs).half()      # tokens > C: H is SPD
    W = (torch.randn(R, C, device=dev, generator=g) * 0.02).half()
    gs = torch.full([C // 128], 128, dtype=torch.int32); bw = torch.full([C // 128], 4, dtype=torch.int32)
    for _ in range(2):
        hook = HessianHook(); hook.add_batch(X); hook.invert(1e-2, True)
        gptq_quant(weight=W, hessian_hook=hook, group_sizes=gs, group_bit_widths=bw, scale_bit_width=None, gptq_use_kernel=False,
                   gptq_block_sizes=gs, quant_symmetric=False, quant_mse=False, save_device=dev)
torch.cuda.synchronize()
print('ok')
