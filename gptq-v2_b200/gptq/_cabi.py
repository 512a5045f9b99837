"""ctypes binding of libgptq_b200.so (C ABI declared in include/gptq_b200.h).

Host-side plumbing only: torch supplies device memory, the current stream and
the device guard; every computation is a call into the sm_100a library.  There
is no CPU fallback: if the library is missing the import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'lib', 'libgptq_b200.so')

# name -> (restype, argtypes); mirrors include/gptq_b200.h one to one
_vp, _i, _f = C.c_void_p, C.c_int, C.c_float
SIGNATURES = {
    'gq_abi_version': (_i, []),
    'gq_error_string': (C.c_char_p, [_i]),
    'gq_launch_count': (C.c_ulonglong, []),
    'gq_accumulate_hessian': (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    'gq_accumulate_hessian_variant': (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    'accumulate_hessian_fp16_fp32_cuda': (_i, [_vp, _vp, _i, _i]),
    'accumulate_hessian_bf16_fp32_cuda': (_i, [_vp, _vp, _i, _i]),
    'gq_invert_workspace_floats': (C.c_size_t, [_i]),
    'gq_hessian_invert': (_i, [_vp, _i, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp]),
    'gq_hessian_invert_dev': (_i, [_vp, _i, _vp, _f, _vp, _f, _vp, _vp, _vp, _vp, _vp]),
    'gq_upper_packed_floats': (C.c_size_t, [_i]),
    'gq_pack_upper': (_i, [_vp, _i, _vp, _vp]),
    'gq_unpack_upper': (_i, [_vp, _i, _vp, _vp]),
    'gq_quantize_range': (_i, [_vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'gq_gptq_solve': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    'gq_gptq_solve_static': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    'gq_gptq_solve_ex': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _f, _i, _i, _i, _vp]),
    'gq_find_params': (_i, [_vp, _i, _i, _i, _vp, _i, _vp, _vp, _vp]),
    'gq_rtn_quantize': (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    'gq_mse_refine': (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _f, _vp, _vp, _vp, _vp]),
    'gq_prep_weight_t': (_i, [_vp, _i, _vp, _vp, _i, _i, _vp]),
    'gq_prep_weight_t2': (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    'gq_prep_weight_t3': (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    'gq_solve_metrics': (_i, [_vp, _vp, _vp, _i, C.c_size_t, _vp, _vp]),
    'gq_finalize_quant': (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    'gq_pack_int4': (_i, [_vp, _vp, _i, _i, _vp]),
    'gq_unpack_int4': (_i, [_vp, _vp, _i, _i, _vp]),
    'gq_transpose_pack_int4': (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    'gq_w4a16_dequant': (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'gq_reconstruct_dense': (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'gq_llama_rms_norm': (_i, [_vp, _vp, _vp, C.c_longlong, _i, C.c_float, _i, _vp]),
    'gq_llama_rope': (_i, [_vp, _vp, _vp, C.c_longlong, _i, _i, _i, _i, _vp]),
    'gq_llama_silu_mul': (_i, [_vp, _vp, _vp, C.c_longlong, _i, _vp]),
    'gq_gemm_f16_nt': (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    'gq_w4a16_workspace_bytes': (C.c_size_t, [_i, _i, _i]),
    'gq_w4a16_gemm': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    'gq_w4a16_gemv_static': (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    'gq_gemm_tf32x3': (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    'gq_set_tf32x3': (_i, [_i]),
    'gq_set_solve_super': (_i, [_i]),
    'gq_set_solve_fused': (_i, [_i]),
    'gq_set_tf32_kchunk': (_i, [_i]),
    'gq_set_syrk_kchunk': (_i, [_i]),
    'gq_set_solve_debug': (_vp, [_vp]),
    'gq_set_w4a16_path': (_i, [_i]),
    'gq_set_pdl': (_i, [_i]),
    'gq_sgemm': (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _f, _i, _vp]),
}

DTYPE_CODE = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f'{LIB_PATH} not found: build it with `python __graft_entry__.py` (or `make -C gptq-v2_b200/csrc`). '
            'gptq_b200 has no CPU or PyTorch fallback.')
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if lib.gq_abi_version() != 1:
        raise ImportError('libgptq_b200.so ABI version mismatch')
    if os.environ.get('GQ_SOLVE_FUSED', '') == '0':
        lib.gq_set_solve_fused(0)
    return lib


lib = _load()


class GptqB200Error(RuntimeError):
    pass


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib.gq_error_string(rc).decode()
        extra = ''
        if rc == 1 and torch.cuda.is_available():
            try:
                torch.cuda.synchronize()
            except Exception as e:  # surfaces the sticky CUDA error text
                extra = f' [{e}]'
        raise GptqB200Error(f'{what}: {msg} (code {rc}){extra}')


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_of(t: torch.Tensor):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def require_cuda(*tensors) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise GptqB200Error('gptq_b200 kernels take CUDA tensors only (there is no CPU path)')
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise GptqB200Error('all tensors must live on the same CUDA device')
    return dev


def require_contig(**named) -> None:
    for k, t in named.items():
        if t is not None and not t.is_contiguous():
            raise GptqB200Error(f'{k} must be contiguous (row-major)')
