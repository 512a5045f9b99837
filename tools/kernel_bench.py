#!/usr/bin/env python
"""Per-kernel timings with CUDA events (no profiler): K1 SYRK, K2 invert, K3+K4 solve, K5 pack, K6 W4A16.
usage: python tools/kernel_bench.py [syrk] [invert] [solve] [w4a16] [sgemm]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'gptq-v2_b200'))
import torch  # noqa: E402
import gptq  # noqa: E402

dev = torch.device('cuda')
try:
    _pk = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    PEAK_HBM, PEAK_BF16 = _pk['hbm_gbs'] * 1e9, _pk['bf16_tflops'] * 1e12
except Exception:
    PEAK_HBM, PEAK_BF16 = 6560.6e9, 1626.9e12


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def bench_syrk():
    for D in (4096, 11008, 5120, 13824):
        T = 32768
        X = torch.randn(T, D, device=dev).half()
        H = torch.zeros(D, D, device=dev)
        for variant in (1, 2):
            for mode in (0, 1):
                t = timeit(lambda: gptq.hessian_syrk(H, X, mode=mode, variant=variant))
                print(json.dumps({'k': 'syrk', 'D': D, 'T': T, 'variant': variant, 'mode': mode, 'ms': t * 1e3,
                                  'tflops_tri': T * D * (D + 1) / t / 1e12}))
        del X, H


def _supers():
    return tuple(int(v) for v in os.environ.get('GQ_SUPERS', '8').split(','))


def bench_invert():
    for D in (4096, 11008):
        X = torch.randn(2 * D, D, device=dev) * torch.exp(torch.rand(D, device=dev) * 3 - 1.5)
        H = (X.t() @ X)
        perm = torch.argsort(H.diagonal(), descending=True)
        damp = 0.01 * float(H.diagonal().mean())
        for sup in _supers():
            old = gptq.set_solve_super(sup)
            t = timeit(lambda: gptq.hessian_invert(H, perm, 1.0, damp), iters=3, warm=1)
            gptq.set_solve_super(old)
            print(json.dumps({'k': 'invert', 'D': D, 'super': sup, 'ms': t * 1e3, 'tflops_2_3_D3': 2 / 3 * D ** 3 / t / 1e12}))
        torch.backends.cuda.matmul.allow_tf32 = False
        Hp = H[perm][:, perm] + damp * torch.eye(D, device=dev)

        def ref():
            L = torch.linalg.cholesky_ex(Hp)[0]
            Hi = torch.cholesky_inverse(L)
            return torch.linalg.cholesky_ex(Hi, upper=True)[0]
        t2 = timeit(ref, iters=3, warm=1)
        print(json.dumps({'k': 'invert_torch_cusolver', 'D': D, 'ms': t2 * 1e3}))
        del X, H, Hp


def bench_solve():
    shapes = ((4096, 4096), (11008, 4096), (4096, 11008))
    if os.environ.get('GQ_SHAPES'):
        shapes = tuple(tuple(int(v) for v in sh.split('x')) for sh in os.environ['GQ_SHAPES'].split(','))
    for R, C in shapes:
        X = torch.randn(2 * C, C, device=dev) * torch.exp(torch.rand(C, device=dev) * 3 - 1.5)
        H = X.t() @ X
        _, U, _ = gptq.hessian_invert(H, None, 1.0, 0.01 * float(H.diagonal().mean()), want_hp=False)
        W = (torch.randn(R, C, device=dev) * 0.02).half()
        Wt0 = gptq.prep_weight_t(W, None)

        def run():
            Wt = Wt0.clone()
            gptq.gptq_solve(Wt, U, 128, 15.0)
        tc = timeit(lambda: Wt0.clone(), iters=3, warm=1)
        for sup in _supers():
            old = gptq.set_solve_super(sup)
            t = timeit(run, iters=3, warm=1)
            gptq.set_solve_super(old)
            print(json.dumps({'k': 'solve', 'R': R, 'C': C, 'super': sup, 'fused': os.environ.get('GQ_SOLVE_FUSED', '1'), 'ms': (t - tc) * 1e3,
                              'trailing_tflops': R * C * C / (t - tc) / 1e12, 'us_per_block': (t - tc) / (C / 128) * 1e6}))
        del X, H, U, W, Wt0


def bench_quantizer():
    """Quantizer.find_params / .mse on the transposed layout, and gptq_quant end to end with each option."""
    from gptq_py import HessianHook, gptq_quant
    for R, C in ((4096, 4096), (11008, 4096)):
        G = C // 128
        Wt = torch.randn(C, R, device=dev) * 0.02
        maxq = torch.full([G], 15.0, device=dev)
        scale = torch.empty(G, R, device=dev)
        zero = torch.empty_like(scale)
        grid = 1. - torch.arange(0., .8, .01, device=dev)
        t = timeit(lambda: gptq.find_params(Wt, 128, maxq, False, scale, zero))
        print(json.dumps({'k': 'find_params', 'R': R, 'C': C, 'us': t * 1e6, 'GBps': 4 * R * C / t / 1e9}))
        s0 = scale.clone()

        def mse():
            scale.copy_(s0)
            gptq.mse_refine(Wt, 128, maxq, grid, 2.4, scale, zero)
        t = timeit(mse)
        print(json.dumps({'k': 'mse_refine', 'R': R, 'C': C, 'grid': int(grid.numel()), 'norm': 2.4, 'ms': t * 1e3,
                          'G_candidates_per_s': R * C * grid.numel() / t / 1e9}))
    R = C = 4096
    X = (torch.randn(8, 2048, C, device=dev) * torch.exp(torch.rand(C, device=dev) * 3 - 1.5)).half()
    hook = HessianHook()
    hook.add_batch(X)
    hook.invert(1e-2, True)
    W = (torch.randn(R, C, device=dev) * 0.02).half()
    gs = torch.full([C // 128], 128, dtype=torch.int32, device=dev)
    bw = torch.full([C // 128], 4, dtype=torch.int32, device=dev)
    for label, kw in (('plain', {}), ('sym', {'quant_symmetric': True}), ('mse', {'quant_mse': True}),
                      ('exl2+mse', {'quant_mse': True, 'scale_bit_width': torch.tensor(4)})):
        for use_kernel in (True, False):
            args = dict(weight=W, hessian_hook=hook, group_sizes=gs, group_bit_widths=bw, gptq_use_kernel=use_kernel,
                        gptq_block_sizes=gs, quant_mse=False, save_device=dev)
            args.update(kw)
            t = timeit(lambda: gptq_quant(**args), iters=3, warm=1)
            print(json.dumps({'k': 'gptq_quant', 'R': R, 'C': C, 'options': label,
                              'branch': 'kernel' if use_kernel else 'fp32', 'ms': t * 1e3}))


def bench_prep():
    """transposing / permuting weight prep: gather form (perm) vs scatter form (perm_inv)"""
    for R, C in ((4096, 4096), (11008, 4096), (4096, 11008)):
        W = (torch.randn(R, C, device=dev) * 0.02).half()
        perm = torch.randperm(C, device=dev)
        pinv = torch.argsort(perm)
        tg = timeit(lambda: gptq.prep_weight_t2(W, perm))
        ts = timeit(lambda: gptq.prep_weight_t2(W, perm, pinv))
        byts = R * C * (2 + 4 + 2)
        print(json.dumps({'k': 'prep_weight', 'R': R, 'C': C, 'gather_us': tg * 1e6, 'scatter_us': ts * 1e6,
                          'scatter_GBps': byts / ts / 1e9}))


def bench_sgemm():
    torch.backends.cuda.matmul.allow_tf32 = False
    for M, N, K, flags in ((3968, 4096, 128, 1), (10880, 4096, 128, 1), (4096, 4096, 128, 3), (4096, 4096, 4096, 0)):
        A = torch.randn((K, M) if flags & 1 else (M, K), device=dev)
        B = torch.randn(K, N, device=dev)
        Cm = torch.zeros(M, N, device=dev)
        t = timeit(lambda: gptq.sgemm(A, B, Cm, M, N, K, A.stride(0), B.stride(0), N, -1.0, 1.0, flags))
        print(json.dumps({'k': 'sgemm', 'M': M, 'N': N, 'K': K, 'flags': flags, 'us': t * 1e6,
                          'tflops': 2.0 * M * N * K / t / 1e12}))


def bench_tf32():
    for M, N, K, up in ((3968, 4096, 128, 0), (3968, 4096, 128, 2), (10880, 4096, 128, 0), (10880, 4096, 128, 2),
                        (3968, 11008, 128, 0), (10880, 10880, 128, 1), (128, 4096, 128, 0), (4096, 4096, 4096, 0)):
        A = torch.randn(K, M, device=dev)
        B = torch.randn(K, N, device=dev)
        Cm = torch.zeros(M, N, device=dev)
        t = timeit(lambda: gptq.gemm_tf32x3(A, B, Cm, M, N, K, M, N, N, -1.0, up))
        frac = 0.5 if up == 1 else 1.0
        print(json.dumps({'k': 'gemm_tf32x3', 'M': M, 'N': N, 'K': K, 'upper': up, 'us': t * 1e6,
                          'tflops': frac * 2.0 * M * N * K / t / 1e12, 'C_GBps_rw': frac * M * N * 8 / t / 1e9}))


def bench_w4a16():
    """Weights rotate through > 4x L2 of distinct buffers; the launches of one rotation are captured in
    a CUDA graph so that the host (ctypes + allocator) is out of the timed region."""
    # BASELINE.json configs[4]: Llama-2-7B and 70B layer shapes (N = out, K = in), incl. the 70B k/v projection
    shapes = ((4096, 4096), (11008, 4096), (4096, 11008), (8192, 8192), (1024, 8192), (28672, 8192), (8192, 28672))
    if os.environ.get('W4_SHAPES'):                       # e.g. W4_SHAPES=28672x8192 W4_MS=1,8 (profiling runs)
        shapes = tuple(tuple(int(v) for v in sh.split('x')) for sh in os.environ['W4_SHAPES'].split(','))
    m_list = (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096)
    if os.environ.get('W4_PATH'):
        gptq.set_w4a16_path(int(os.environ['W4_PATH']))
    if os.environ.get('W4_MS'):
        m_list = tuple(int(v) for v in os.environ['W4_MS'].split(','))
    for N, K in shapes:
        G = K // 128
        nbuf = max(4, int(600e6 // (N * K // 2)) + 1)
        qs = [torch.randint(-2 ** 31, 2 ** 31 - 1, (N, K // 8), device=dev, dtype=torch.int32) for _ in range(nbuf)]
        sc = (torch.rand(G, N, device=dev) * 0.01 + 1e-3).half()
        qz = torch.randint(0, 16, (G, N), device=dev, dtype=torch.uint8)
        perm = torch.randperm(K, device=dev).to(torch.int32)
        dense = [(torch.randn(N, K, device=dev) * 0.02).half() for _ in range(max(2, min(nbuf, int(2e9 // (N * K * 2)))))]
        for variant, Ms in ((2, m_list),):
            for M in Ms:
                x = torch.randn(M, K, device=dev).half()
                gptq.w4a16_gemm(x, qs[0], sc, qz, perm, 128)
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    for q in qs:
                        gptq.w4a16_gemm(x, q, sc, qz, perm, 128)
                t = timeit(graph.replay, iters=7, warm=2) / nbuf
                del graph
                # a8: the reference evaluates through a dense 16-bit reconstruction + cuBLAS (quant.py:262-279,
                # llama.py:172-177); the GEMM alone (reconstruction not timed), dense weights rotated the same way
                g2 = torch.cuda.CUDAGraph()
                torch.matmul(x, dense[0].t())
                torch.cuda.synchronize()
                with torch.cuda.graph(g2):
                    for wd in dense:
                        torch.matmul(x, wd.t())
                td = timeit(g2.replay, iters=7, warm=2) / len(dense)
                del g2
                byts = N * K / 2 + N * G * 3 + 2 * M * K + 2 * M * N
                print(json.dumps({'k': 'w4a16', 'variant': variant, 'M': M, 'N': N, 'K': K, 'us': t * 1e6,
                                  'GBps': byts / t / 1e9, 'tflops': 2.0 * M * N * K / t / 1e12,
                                  'frac_hbm': byts / t / PEAK_HBM, 'frac_bf16_peak': 2.0 * M * N * K / t / PEAK_BF16,
                                  'dense_cublas_us': td * 1e6, 'speedup_vs_dense_cublas': td / t}))
        del qs, dense


def bench_gemmnt():
    """gq_gemm_f16_nt (the dense GEMM of K6's large-batch route) beside torch.matmul (cuBLAS) on the same operands."""
    for N, K in ((4096, 4096), (11008, 4096), (4096, 11008), (8192, 8192), (28672, 8192), (8192, 28672)):
        w = (torch.randn(N, K, device=dev) * 0.02).half()
        for M in (128, 256, 512, 1024, 2048, 4096):
            x = torch.randn(M, K, device=dev).half()
            t = timeit(lambda: gptq.gemm_f16_nt(x, w), iters=7, warm=2)
            tb = timeit(lambda: torch.matmul(x, w.t()), iters=7, warm=2)
            print(json.dumps({'k': 'gemm_f16_nt', 'M': M, 'N': N, 'K': K, 'us': t * 1e6, 'tflops': 2.0 * M * N * K / t / 1e12,
                              'frac_bf16_peak': 2.0 * M * N * K / t / PEAK_BF16, 'cublas_us': tb * 1e6, 'vs_cublas': tb / t}))
        del w


def bench_refext():
    """The reference's own CUDA extension (oracle/_ref/gptq_c.so, built from /root/reference/package/gptq) on this
    B200, beside ours: a2 CUTLASS Hessian GEMM and a6 gptq_quantize_range over a whole 4096x4096 linear."""
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import ref_ext
    ref = ref_ext.load()
    if ref is None:
        print(json.dumps({'k': 'refext', 'unavailable': 'oracle/_ref/gptq_c.so not built'}))
        return
    torch.backends.cuda.matmul.allow_tf32 = False
    for D in (4096, 11008):
        T = 32768
        X = torch.randn(T, D, device=dev).half()
        H = torch.zeros(D, D, device=dev)
        t_ref = timeit(lambda: ref.accumulate_hessian_fp16_fp32(H, X))
        t_our2 = timeit(lambda: gptq.accumulate_hessian(H, X))
        t_our0 = timeit(lambda: gptq.hessian_syrk(H, X, mode=0))
        print(json.dumps({'k': 'refext_hessian', 'D': D, 'T': T, 'ref_cutlass_ms': t_ref * 1e3,
                          'ref_gemm_tflops': 2.0 * T * D * D / t_ref / 1e12, 'ours_mode2_ms': t_our2 * 1e3,
                          'ours_mode0_ms': t_our0 * 1e3, 'speedup_mode0': t_ref / t_our0}))
        del X, H
    R = C = 4096
    Xc = torch.randn(2 * C, C, device=dev) * torch.exp(torch.rand(C, device=dev) * 3 - 1.5)
    Hc = Xc.t() @ Xc
    _, U, _ = gptq.hessian_invert(Hc, None, 1.0, 0.01 * float(Hc.diagonal().mean()), want_hp=False)
    W = (torch.randn(R, C, device=dev) * 0.02)
    Wt0 = W.t().contiguous()
    w3 = Wt0.view(C // 128, 128, R)
    xmax, xmin = w3.amax(dim=1).relu(), -(-w3.amin(dim=1)).relu()
    scale = (xmax - xmin) / 15.0 + 1e-12
    qzero = (-xmin / scale).round()
    quant_t, err_t = torch.empty_like(Wt0), torch.empty_like(Wt0)
    q_t = torch.empty_like(Wt0, dtype=torch.int16)

    def run(fn):
        wt = Wt0.clone()
        for g in range(C // 128):
            fn(quant_t, scale[g], q_t, qzero[g], 15.0, U, wt, err_t, g * 128, (g + 1) * 128)
    t_ref = timeit(lambda: run(ref.gptq_quantize_range), iters=3, warm=1)
    t_our = timeit(lambda: run(gptq.gptq_quantize_range), iters=3, warm=1)
    print(json.dumps({'k': 'refext_quantize_range', 'R': R, 'C': C, 'ref_ms': t_ref * 1e3, 'ours_ms': t_our * 1e3,
                      'speedup': t_ref / t_our}))


if __name__ == '__main__':
    if os.environ.get('GQ_TF32_MODE'):
        gptq.set_tf32x3(int(os.environ['GQ_TF32_MODE']))
    which = sys.argv[1:] or ['syrk', 'invert', 'solve', 'sgemm', 'tf32', 'w4a16', 'refext']
    for w in which:
        globals()['bench_' + w]()
