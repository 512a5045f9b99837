"""Loader for tests/golden/*.npz (written by oracle/gen_golden.py from the real reference)."""
import glob
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
_DT = {'float16': torch.float16, 'bfloat16': torch.bfloat16}


def names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, 'r*.npz')))


def qnames():
    """The Quantizer-option cases."""
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, 'q_*.npz')))


def loadq(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    d = {k: z[k] for k in z.files}
    dt = _DT[str(d['dtype'])]
    sb = int(d['scale_bits'])
    out = dict(name=name, R=int(d['R']), C=int(d['C']), bits=[int(b) for b in d['bits']], branch=str(d['branch']),
               sym=bool(int(d['sym'])), scale_bits=(None if sb < 0 else sb), mse=bool(int(d['mse'])),
               act_order=bool(int(d['act_order'])), dtype=dt)
    out['W'] = torch.from_numpy(d['W_bits'].copy()).view(dt)
    for k in ('scale', 'W_hat', 'qweight', 'qzero', 'perm_inv', 'group_bit_widths', 'U', 'qscale', 'sscale'):
        out[k] = torch.from_numpy(d[k].copy()) if k in d else None
    out['perm'] = torch.from_numpy(d['perm'].copy()).to(torch.int64) if ('perm' in d and out['act_order']) else None
    out['metrics'] = dict(zip(('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'), d['metrics'].tolist()))
    return out


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    d = {k: z[k] for k in z.files}
    dt = _DT[str(d['dtype'])]
    out = dict(name=name, R=int(d['R']), C=int(d['C']), n_seq=int(d['n_seq']), seqlen=int(d['seqlen']),
               bits=int(d['bits']), act_order=bool(int(d['act_order'])), dtype=dt)
    out['W'] = torch.from_numpy(d['W_bits'].copy()).view(dt)
    out['X'] = torch.from_numpy(d['X_bits'].copy()).view(dt)
    for k in ('H_raw', 'H_damped', 'U', 'scale', 'W_hat'):
        out[k] = torch.from_numpy(d[k].copy())
    out['perm'] = torch.from_numpy(d['perm'].copy()).to(torch.int64)
    out['qweight'] = torch.from_numpy(d['qweight'].copy())
    out['qzero'] = torch.from_numpy(d['qzero'].copy())
    out['perm_inv'] = torch.from_numpy(d['perm_inv'].copy())
    out['metrics'] = dict(zip(('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'), d['metrics'].tolist()))
    return out


def bnames():
    """GPTQ blocks != quantisation groups, non-uniform group sizes (fp32 branch of the reference)."""
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, 'b_*.npz')))


def loadb(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    d = {k: z[k] for k in z.files}
    dt = _DT[str(d['dtype'])]
    sb = int(d['scale_bits']) if 'scale_bits' in d else -1
    out = dict(name=name, R=int(d['R']), C=int(d['C']), group_sizes=[int(v) for v in d['group_sizes']],
               bits=[int(v) for v in d['bits']], block_sizes=[int(v) for v in d['block_sizes']],
               act_order=bool(int(d['act_order'])), dtype=dt, rtn=bool(int(d['rtn'])) if 'rtn' in d else False,
               sym=bool(int(d['sym'])) if 'sym' in d else False, mse=bool(int(d['mse'])) if 'mse' in d else False,
               scale_bits=(None if sb < 0 else sb))
    out['W'] = torch.from_numpy(d['W_bits'].copy()).view(dt)
    for k in ('U', 'scale', 'W_hat', 'qweight', 'qzero', 'perm_inv', 'qscale', 'sscale', 'group_sizes_out',
              'group_bit_widths_out'):
        out[k] = torch.from_numpy(d[k].copy()) if k in d else None
    if out['rtn']:
        out['U'] = None
    out['perm'] = torch.from_numpy(d['perm'].copy()).to(torch.int64) if out['act_order'] else None
    out['metrics'] = dict(zip(('gptq_error', 'gptq_norm', 'l2.0_error', 'l2.0_norm'), d['metrics'].tolist()))
    return out
