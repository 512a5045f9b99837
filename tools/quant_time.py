#!/usr/bin/env python
"""K3+K4+K5 in isolation: one gptq_quant call per Llama-2-7B linear shape (4-bit g128 act-order, the bench's settings) timed with
CUDA events, whole call (deferred metrics: no host sync inside) and by part (prep / solve / metrics / transpose+pack / rest).
    python tools/quant_time.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402


def timed(fn, reps=5):
    ts = []
    out = None
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts[1:])[len(ts[1:]) // 2], out


def main():
    import gptq
    from gptq_py import HessianHook, gptq_quant
    dev = torch.device('cuda:0')
    hooks = {}
    shapes = [tuple(int(v) for v in a.split('x')) for a in sys.argv[1:]] or [(4096, 4096), (11008, 4096), (4096, 11008)]
    for R, C in shapes:
        if C not in hooks:
            g = torch.Generator(device=dev).manual_seed(C)
            cs = torch.exp(torch.rand(C, device=dev, generator=g) * 3.0 - 1.5)
            X = (torch.randn(16, 2048, C, device=dev, generator=g) * cs).half()
            h = HessianHook()
            h.add_batch(X)
            del X
            h.invert(1e-2, True)
            hooks[C] = h
        hook = hooks[C]
        g = torch.Generator(device=dev).manual_seed(R + C)
        W = (torch.randn(R, C, device=dev, generator=g) * 0.02).half()
        gs = torch.full([C // 128], 128, dtype=torch.int32)
        bw = torch.full([C // 128], 4, dtype=torch.int32)
        kw = dict(group_sizes=gs, group_bit_widths=bw, scale_bit_width=None, gptq_use_kernel=False, gptq_block_sizes=gs,
                  quant_symmetric=False, quant_mse=False, save_device=dev)
        whole, _ = timed(lambda: gptq_quant(weight=W, hessian_hook=hook, defer_metrics=True, **kw))
        sync_call, _ = timed(lambda: gptq_quant(weight=W, hessian_hook=hook, **kw))
        U = hook.hessian_inv
        prep, (Wt, orig_t, colsq) = timed(lambda: gptq.prep_weight_t2(W, hook.perm, hook.perm_inv))
        scale_gr = torch.empty(C // 128, R, device=dev)
        qzero_gr = torch.empty_like(scale_gr)

        def solve():
            Wt2 = Wt.clone()
            return gptq.gptq_solve_ex(Wt2, U, scale_gr, qzero_gr, 128, [15.0] * (C // 128), mode=0)
        clone_t, _ = timed(lambda: Wt.clone())
        solve_t, (quant_t, codes_t, err_t) = timed(solve)
        met, _ = timed(lambda: gptq.solve_metrics(err_t, quant_t, orig_t))
        tp, _ = timed(lambda: gptq.transpose_pack_int4(codes_t, want_codes=True, want_packed=False))
        glue, _ = timed(lambda: (scale_gr.t().to(torch.float16).contiguous(), qzero_gr.t().to(torch.uint8).contiguous(),
                                 (W.abs() ** 2.).mean()))
        extra = {}
        if (R, C) == (4096, 4096):
            from gptq_py import gptq_quant_stacked
            W3 = [W, W.clone(), W.clone()]
            t3, _ = timed(lambda: gptq_quant_stacked(W3, hook, defer_metrics=True, **kw))
            extra = {'stacked_3x4096_rows_ms': round(t3, 3), 'three_separate_calls_ms': round(3 * whole, 3)}
        print(json.dumps({**extra, 'R': R, 'C': C, 'gptq_quant_ms': round(whole, 3), 'gptq_quant_with_host_sync_ms': round(sync_call, 3),
                          'prep_ms': round(prep, 3), 'solve_ms': round(solve_t - clone_t, 3), 'metrics_ms': round(met, 3),
                          'transpose_ms': round(tp, 3), 'scale_zero_l2norm_glue_ms': round(glue, 3)}))


if __name__ == '__main__':
    main()
