"""Freeze outputs of the REAL reference (imported from /root/reference) as small
golden fixtures under tests/golden/.  Runs only in the build container -- the
GPU box has no /root/reference; the tests read the committed .npz files.

    python oracle/gen_golden.py            # (re)writes tests/golden/*.npz

What is frozen, per case (inputs are regenerated from the seed by
oracle.gptq_oracle.synth_linear, and stored as raw 16-bit patterns (W_bits,
X_bits) so the fixture is self-contained):
  * HessianHook.add_batch(use_kernel=False) -> H            (gptq_py.py:19-33)
  * HessianHook.invert(1e-2, act_order)     -> U, perm, damped H  (:39-66)
  * gptq_quant(gptq_use_kernel=False, quant_mse=False)      (:158-251)
        -> quant_meta (qweight/scale/qzero/perm_inv) + 4 metrics
  * quant.reconstruct_nn_linear(...).weight                 (quant.py:262-279)
The reference's gptq_use_kernel=True branch needs its CUDA extension and
cannot run here; its semantics are pinned by the C++/CUDA source restatement
in oracle.gptq_oracle.quantize_range_ref only ("parity unpinned" for that
compat mode beyond self-consistency).

`python oracle/gen_golden.py quantizer` writes only the q_*.npz cases: symmetric grids, MSE grid search, EXL2 quantised
scales and per-group bit widths (quant.py:40-69,99-112,120-165) through the reference's round-to-nearest and fp32
branches (gptq_py.py:101-124,158-193).
`python oracle/gen_golden.py blocks` writes only the b_*.npz cases: GPTQ blocks larger / smaller than the quantisation
groups and non-uniform group sizes through the reference's fp32 branch (gptq_py.py:158-193).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = '/root/reference'
sys.path.insert(0, ROOT)

CASES = [
    # name,        R,   C,  n_seq, seqlen, bits, act_order, dtype
    ('r64_c256_b4', 64, 256, 8, 96, 4, True, torch.float16),
    ('r96_c384_b3', 96, 384, 6, 128, 3, True, torch.float16),
    ('r128_c512_b4_noact', 128, 512, 4, 256, 4, False, torch.float16),
    ('r80_c256_b4_bf16', 80, 256, 8, 64, 4, True, torch.bfloat16),
]


def import_reference():
    """gptq_py.py:3 imports the CUDA extension package `gptq` unconditionally;
    give it a two-function stub (never called on the fp32 path)."""
    stub = types.ModuleType('gptq')

    def _nope(*a, **k):
        raise RuntimeError('reference CUDA extension is not available on CPU')
    stub.accumulate_hessian = _nope
    stub.gptq_quantize_range = _nope
    sys.modules['gptq'] = stub
    sys.path.insert(0, REF)
    import gptq_py  # noqa
    import quant  # noqa
    return gptq_py, quant


def main():
    from oracle import gptq_oracle as O
    gptq_py, quant = import_reference()
    torch.set_num_threads(1)   # LAPACK/BLAS summation order fixed for the freeze
    out_dir = os.path.join(ROOT, 'tests', 'golden')
    os.makedirs(out_dir, exist_ok=True)
    for name, R, C, n_seq, seqlen, bits, act, dtype in CASES:
        W, X = O.synth_linear(R, C, n_seq, seqlen, dtype=dtype, seed=100 + R + C)
        hook = gptq_py.HessianHook()
        for i in range(0, n_seq, 2):
            hook.add_batch(X[i:i + 2], use_kernel=False)
        H_raw = hook.hessian.clone()
        hook.invert(damp_ratio=1e-2, act_order=act)
        G = C // 128
        gs = torch.full([G], 128, dtype=torch.int32)
        bw = torch.full([G], bits, dtype=torch.int32)
        res = gptq_py.gptq_quant(weight=W, hessian_hook=hook, group_sizes=gs, group_bit_widths=bw,
                                 scale_bit_width=None, gptq_use_kernel=False, gptq_block_sizes=gs,
                                 quant_symmetric=False, quant_mse=False)
        lin = quant.reconstruct_nn_linear(res['quant_meta'], dtype=dtype)
        m = res['quant_meta']
        to_np = lambda t: t.to(torch.float32).numpy() if t.dtype in (torch.float16, torch.bfloat16) else t.numpy()
        np.savez_compressed(
            os.path.join(out_dir, name + '.npz'),
            R=R, C=C, n_seq=n_seq, seqlen=seqlen, bits=bits, act_order=int(act),
            dtype=str(dtype).split('.')[-1],
            W_bits=W.view(torch.int16).numpy(), X_bits=X.view(torch.int16).numpy(),
            H_raw=H_raw.numpy(), H_damped=hook.hessian.numpy(), U=hook.hessian_inv.numpy(),
            perm=(hook.perm.numpy() if hook.perm is not None else np.arange(C)),
            qweight=m['qweight'].numpy(), scale=to_np(m['scale']), qzero=m['qzero'].numpy(),
            perm_inv=m['perm_inv'].numpy(),
            metrics=np.array([res['metrics'][k] for k in ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')],
                             dtype=np.float64),
            W_hat=to_np(lin.weight.data),
        )
        print('wrote', name, {k: f'{v:.4e}' for k, v in res['metrics'].items()})


# ---- the remaining Quantizer options, through the reference's own RTN / fp32 branches
Q_CASES = [
    # name,            R,  C, bits per group,  branch, sym,  scale_bits, mse, act_order, dtype
    ('q_sym_b4',       64, 256, [4, 4],        'pure', True,  None, False, True, torch.float16),
    ('q_mse_b4',       64, 256, [4, 4],        'pure', False, None, True,  True, torch.float16),
    ('q_mse_sym_b3',   48, 256, [3, 3],        'pure', True,  None, True,  False, torch.float16),
    ('q_exl2_b4',      64, 256, [4, 4],        'pure', False, 4,    False, True, torch.float16),
    ('q_exl2_mse_mix', 64, 384, [3, 4, 2],     'pure', False, 4,    True,  True, torch.float16),
    ('q_mixed_bits',   64, 384, [2, 8, 4],     'pure', False, None, False, True, torch.bfloat16),
    ('q_rtn_mse_sym',  64, 256, [4, 3],        'rtn',  True,  None, True,  False, torch.float16),
    ('q_rtn_exl2_mse', 32, 256, [4, 4],        'rtn',  False, 6,    True,  False, torch.bfloat16),
]


def main_quantizer():
    from oracle import gptq_oracle as O
    gptq_py, quant = import_reference()
    torch.set_num_threads(1)
    out_dir = os.path.join(ROOT, 'tests', 'golden')
    for name, R, C, bits, branch, sym, sbits, mse, act, dtype in Q_CASES:
        n_seq, seqlen = 4, 128
        W, X = O.synth_linear(R, C, n_seq, seqlen, dtype=dtype, seed=500 + R + C + len(name))
        hook = None
        if branch != 'rtn':
            hook = gptq_py.HessianHook()
            for i in range(0, n_seq, 2):
                hook.add_batch(X[i:i + 2], use_kernel=False)
            hook.invert(damp_ratio=1e-2, act_order=act)
        G = C // 128
        gs = torch.full([G], 128, dtype=torch.int32)
        bw = torch.tensor(bits, dtype=torch.int32)
        res = gptq_py.gptq_quant(weight=W, hessian_hook=hook, group_sizes=gs, group_bit_widths=bw,
                                 scale_bit_width=(torch.tensor(sbits) if sbits is not None else None),
                                 gptq_use_kernel=False, gptq_block_sizes=gs, quant_symmetric=sym, quant_mse=mse,
                                 quant_max_shrink=.8, quant_n_grid=100, quant_norm=2.4, quant_use_kernel=False)
        lin = quant.reconstruct_nn_linear(res['quant_meta'], dtype=dtype)
        m = res['quant_meta']
        to_np = lambda t: t.to(torch.float32).numpy() if t.dtype in (torch.float16, torch.bfloat16) else t.numpy()
        extra = {}
        if m['sscale'] is not None:
            extra = dict(qscale=m['qscale'].numpy(), sscale=to_np(m['sscale']), scale_bit_width=m['scale_bit_width'].numpy())
        if hook is not None:
            extra.update(U=hook.hessian_inv.numpy(), perm=(hook.perm.numpy() if hook.perm is not None else np.arange(C)))
        np.savez_compressed(
            os.path.join(out_dir, name + '.npz'),
            R=R, C=C, bits=np.array(bits), branch=branch, sym=int(sym), scale_bits=(-1 if sbits is None else sbits),
            mse=int(mse), act_order=int(act), dtype=str(dtype).split('.')[-1],
            W_bits=W.view(torch.int16).numpy(),
            qweight=m['qweight'].numpy(), scale=to_np(m['scale']), qzero=m['qzero'].numpy(), perm_inv=m['perm_inv'].numpy(),
            group_bit_widths=m['group_bit_widths'].numpy(),
            metrics=np.array([res['metrics'][k] for k in ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')],
                             dtype=np.float64),
            W_hat=to_np(lin.weight.data), **extra)
        print('wrote', name, {k: f'{v:.4e}' for k, v in res['metrics'].items()})


# ---- GPTQ blocks that differ from the quantisation groups, non-uniform group sizes (gptq_py.py:101-124, 158-193)
B_CASES = [
    # name,                R,  C,  group sizes,          bits,            block sizes (None = round to nearest), act_order, dtype, options
    ('b_blk256_g128',      64, 512, [128] * 4,           [4] * 4,         [256, 256],             True,  torch.float16, {}),
    ('b_blk64_g128',       64, 256, [128, 128],          [4, 4],          [64] * 4,               True,  torch.float16, {}),
    ('b_blk128_g32',       48, 256, [32] * 8,            [3] * 8,         [128, 128],             False, torch.bfloat16, {}),
    ('b_groups_64_128_64', 64, 256, [64, 128, 64],       [4, 3, 4],       [128, 128],             True,  torch.float16, {}),
    # round 2: the MSE search over the BLOCK's columns (:178-184), EXL2 scales, sizes that do not nest, any sizes on the
    # round-to-nearest branch, EXL2 group ordering by (bit width, group size)
    ('b_blk256_g128_mse',  64, 512, [128] * 4,           [4, 3, 4, 4],    [256, 256],             True,  torch.float16, {'mse': True}),
    ('b_blk64_g128_exl2',  48, 256, [128, 128],          [4, 4],          [64] * 4,               True,  torch.float16, {'scale_bits': 4, 'mse': True}),
    ('b_nonnest_96_160',   64, 256, [96, 160],           [4, 4],          [128, 128],             True,  torch.float16, {}),
    ('b_groups_sym_mse',   48, 320, [64, 128, 32, 96],   [4, 3, 4, 2],    [64, 128, 32, 96],      True,  torch.float16, {'sym': True, 'mse': True}),
    ('b_rtn_48_80_128',    40, 256, [48, 80, 128],       [4, 3, 4],       None,                   False, torch.float16, {'mse': True}),
    ('b_rtn_exl2_order',   40, 320, [64, 128, 64, 64],   [3, 4, 4, 3],    None,                   False, torch.bfloat16, {'scale_bits': 4}),
]


def main_blocks(only_new: bool = False):
    from oracle import gptq_oracle as O
    gptq_py, quant = import_reference()
    torch.set_num_threads(1)
    out_dir = os.path.join(ROOT, 'tests', 'golden')
    for name, R, C, gsz, bits, bsz, act, dtype, opt in B_CASES:
        if only_new and os.path.exists(os.path.join(out_dir, name + '.npz')):
            continue
        n_seq, seqlen = 4, 128
        W, X = O.synth_linear(R, C, n_seq, seqlen, dtype=dtype, seed=900 + R + C + len(name))
        sbits = opt.get('scale_bits')
        kw = dict(group_sizes=torch.tensor(gsz, dtype=torch.int32), group_bit_widths=torch.tensor(bits, dtype=torch.int32),
                  scale_bit_width=(torch.tensor(sbits, dtype=torch.int32) if sbits is not None else None),
                  gptq_use_kernel=False, quant_symmetric=bool(opt.get('sym', False)), quant_mse=bool(opt.get('mse', False)))
        if bsz is None:
            hook = None
            res = gptq_py.gptq_quant(weight=W, hessian_hook=None, **kw)
        else:
            hook = gptq_py.HessianHook()
            for i in range(0, n_seq, 2):
                hook.add_batch(X[i:i + 2], use_kernel=False)
            hook.invert(damp_ratio=1e-2, act_order=act)
            res = gptq_py.gptq_quant(weight=W, hessian_hook=hook, gptq_block_sizes=torch.tensor(bsz, dtype=torch.int32), **kw)
        lin = quant.reconstruct_nn_linear(res['quant_meta'], dtype=dtype)
        m = res['quant_meta']
        to_np = lambda t: t.to(torch.float32).numpy() if t.dtype in (torch.float16, torch.bfloat16) else t.numpy()
        extra = {}
        if m['qscale'] is not None:
            extra = dict(qscale=m['qscale'].numpy(), sscale=to_np(m['sscale']))
        np.savez_compressed(
            os.path.join(out_dir, name + '.npz'),
            R=R, C=C, group_sizes=np.array(gsz), bits=np.array(bits), block_sizes=np.array(bsz if bsz is not None else []),
            rtn=int(bsz is None), act_order=int(act and hook is not None),
            sym=int(bool(opt.get('sym', False))), mse=int(bool(opt.get('mse', False))), scale_bits=(-1 if sbits is None else sbits),
            dtype=str(dtype).split('.')[-1], W_bits=W.view(torch.int16).numpy(),
            U=(hook.hessian_inv.numpy() if hook is not None else np.zeros((0, 0), dtype=np.float32)),
            perm=(hook.perm.numpy() if hook is not None and hook.perm is not None else np.arange(C)),
            qweight=m['qweight'].numpy(), scale=to_np(m['scale']), qzero=m['qzero'].numpy(), perm_inv=m['perm_inv'].numpy(),
            group_sizes_out=m['group_sizes'].numpy(), group_bit_widths_out=m['group_bit_widths'].numpy(),
            metrics=np.array([res['metrics'][k] for k in ('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm')],
                             dtype=np.float64),
            W_hat=to_np(lin.weight.data), **extra)
        print('wrote', name, {k: f'{v:.4e}' for k, v in res['metrics'].items()})


if __name__ == '__main__':
    if len(sys.argv) > 1 and sys.argv[1] == 'quantizer':
        main_quantizer()
    elif len(sys.argv) > 1 and sys.argv[1] == 'blocks':
        main_blocks()
    elif len(sys.argv) > 1 and sys.argv[1] == 'blocks-new':
        main_blocks(only_new=True)                     # add cases without touching the fixtures already committed
    else:
        main()
        main_quantizer()
        main_blocks()
