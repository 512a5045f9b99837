import os, sys
ROOT = os.environ.get('GRAFT_REPO_ROOT', '/root/repo')
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + '/gptq-v2_b200')
import torch, gptq
dev = 'cuda'
g = torch.Generator(device=dev).manual_seed(1)
for (M, N, K) in ((4992, 1708, 128), (4992, 1708, 1024), (3968, 1368, 1024), (13696, 1728, 1024), (4992, 2560, 1024)):
    A = torch.randn(K, M, device=dev, generator=g); B = torch.randn(K, N, device=dev, generator=g)
    C0 = torch.randn(M, N, device=dev, generator=g)
    want = C0.double() - A.double().t() @ B.double()
    for mode in (2, 3):
        prev = gptq.set_tf32x3(mode)
        C = C0.clone()
        gptq.gemm_tf32x3(A, B, C, M, N, K, M, N, N, -1.0, False)
        torch.cuda.synchronize()
        gptq.set_tf32x3(prev)
        print('gemm', M, N, K, 'mode', mode, 'max err', float((C.double() - want).abs().max()))
for D in (5120, 11008):
    X = torch.randn(2 * D, D, device=dev, generator=g) * torch.exp(torch.rand(D, device=dev, generator=g) * 3 - 1.5)
    H = X.t() @ X
    perm = torch.argsort(H.diagonal(), descending=True)
    damp = 0.01 * float(H.diagonal().mean())
    outs = {}
    for mode in (2, 3):
        prev = gptq.set_tf32x3(mode)
        _, U, info = gptq.hessian_invert(H, perm, 1.0, damp)
        torch.cuda.synchronize()
        gptq.set_tf32x3(prev)
        outs[mode] = U
        print('invert', D, 'mode', mode, 'info', int(info), 'finite', bool(torch.isfinite(U).all()))
    print('invert', D, 'rel diff 3 vs 2', float((outs[3] - outs[2]).norm() / outs[2].norm()))
    del X, H, outs
