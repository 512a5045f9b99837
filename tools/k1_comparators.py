#!/usr/bin/env python
"""The two comparators SURVEY section 2a names for K1, timed on the same box and shapes as gq_accumulate_hessian:

  triton   a Triton kernel of the reference's design (/root/reference/gptq_triton/accumulate_hessian.py:55-149, the live call site
           of /root/reference/gptq_py.py:36): tl.dot tiles of X^T X with fp32 accumulators, blocks above the diagonal skipped,
           every computed block added to H and mirrored.  Written here from that description (the reference tree does not
           exist on the GPU box); the best of the tile shapes the reference autotunes over is reported.
  addmm    the reference's fall-back, cuBLAS in fp32 with TF32 off: H.addmm_(X.float().t(), X.float()) (gptq_py.py:31-33),
           with and without the fp16 -> fp32 conversion of X inside the timed region.

    python tools/k1_comparators.py [D ...]          (T = 32768 tokens per call, the bench's launch)
Comparator only: Triton is not used anywhere in the product."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import triton  # noqa: E402
import triton.language as tl  # noqa: E402


@triton.jit
def xtx_lower_kernel(h_ptr, x_ptr, D, T, BM: tl.constexpr, BN: tl.constexpr, BK: tl.constexpr, GROUP: tl.constexpr):
    pid = tl.program_id(0)
    nm = tl.cdiv(D, BM)
    nn = tl.cdiv(D, BN)
    per_group = GROUP * nn
    gid = pid // per_group
    m_first = gid * GROUP
    gsz = min(nm - m_first, GROUP)
    pm = m_first + (pid % per_group) % gsz
    pn = (pid % per_group) // gsz
    if (pm + 1) * BM <= pn * BN:                      # block strictly above the diagonal: its mirror image computes it
        return
    rm = (pm * BM + tl.arange(0, BM)) % D
    rn = (pn * BN + tl.arange(0, BN)) % D
    rk = tl.arange(0, BK)
    a_ptrs = x_ptr + rm[:, None] + rk[None, :] * D    # X^T: (D, T) view of the (T, D) row-major input
    b_ptrs = x_ptr + rk[:, None] * D + rn[None, :]
    acc = tl.zeros((BM, BN), dtype=tl.float32)
    for k in range(0, tl.cdiv(T, BK)):
        a = tl.load(a_ptrs, mask=rk[None, :] < T - k * BK, other=0.)
        b = tl.load(b_ptrs, mask=rk[:, None] < T - k * BK, other=0.)
        acc = tl.dot(a, b, acc)
        a_ptrs += BK * D
        b_ptrs += BK * D
    om = pm * BM + tl.arange(0, BM)
    on = pn * BN + tl.arange(0, BN)
    mask = (om[:, None] < D) & (on[None, :] < D)
    c_ptrs = h_ptr + om[:, None] * D + on[None, :]
    acc += tl.load(c_ptrs, mask=mask)
    tl.store(c_ptrs, acc, mask=mask)
    if pm * BM >= (pn + 1) * BN:                      # strictly below the diagonal: mirror into the upper triangle
        tl.store(h_ptr + om[:, None] + on[None, :] * D, acc, mask=mask)


CONFIGS = [(128, 256, 64, 8, 8, 3), (128, 128, 32, 8, 4, 4), (64, 256, 32, 8, 4, 4), (128, 64, 32, 8, 4, 4), (64, 128, 32, 8, 4, 4),
           (128, 128, 64, 8, 8, 3), (256, 128, 64, 8, 8, 3)]


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    import gptq
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = torch.device('cuda:0')
    T = 32768
    for D in [int(a) for a in sys.argv[1:]] or [4096, 11008]:
        X = torch.randn(T, D, device=dev).half()
        H64 = None
        out = {'D': D, 'T': T, 'flops_triangular': T * D * (D + 1)}
        H = torch.zeros(D, D, device=dev)
        ms = timeit(lambda: gptq.hessian_syrk(H, X, mode=0))
        out['this_build_K1'] = {'ms': round(ms, 3), 'tflops_tri': round(T * D * (D + 1) / ms / 1e9, 1)}
        H.zero_(); gptq.hessian_syrk(H, X, mode=0); torch.cuda.synchronize()
        Hk = torch.triu(H).clone()
        best = None
        for BM, BN, BK, G, warps, stages in CONFIGS:
            try:
                grid = (triton.cdiv(D, BM) * triton.cdiv(D, BN),)
                Ht = torch.zeros(D, D, device=dev)
                fn = lambda: xtx_lower_kernel[grid](Ht, X, D, T, BM=BM, BN=BN, BK=BK, GROUP=G, num_warps=warps, num_stages=stages)
                ms_t = timeit(fn, iters=3, warm=1)
                Ht.zero_(); fn(); torch.cuda.synchronize()
                rel = float((torch.triu(Ht) - Hk).norm() / Hk.norm())
                rec = {'tile': [BM, BN, BK], 'warps': warps, 'stages': stages, 'ms': round(ms_t, 3),
                       'tflops_tri': round(T * D * (D + 1) / ms_t / 1e9, 1), 'rel_fro_vs_K1': rel}
                if best is None or ms_t < best['ms']:
                    best = rec
            except Exception as e:                                         # a tile shape that does not fit: skip it
                out.setdefault('triton_skipped', []).append([BM, BN, BK, type(e).__name__])
        out['triton_best'] = best
        Hf = torch.zeros(D, D, device=dev)
        Xf = X.float()
        ms_a = timeit(lambda: Hf.addmm_(Xf.t(), Xf), iters=3, warm=1)
        out['cublas_fp32_addmm'] = {'ms': round(ms_a, 3), 'tflops_full': round(2 * T * D * D / ms_a / 1e9, 1)}
        ms_b = timeit(lambda: Hf.addmm_(X.float().t(), X.float()), iters=3, warm=1)
        out['cublas_fp32_addmm_with_conversion'] = {'ms': round(ms_b, 3)}
        out['speedup_vs_triton'] = round(best['ms'] / ms, 2) if best else None
        out['speedup_vs_addmm'] = round(ms_b / ms, 2)
        print(json.dumps(out), flush=True)
        del X, Xf, H, Hf


if __name__ == '__main__':
    main()
