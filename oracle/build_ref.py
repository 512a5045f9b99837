"""Compile the REFERENCE's own CUDA extension `gptq_c` into oracle/_ref/  --  TEST INFRASTRUCTURE ONLY.

Sources are compiled where they lie under /root/reference/package/gptq (pybind.cpp, accumulate_hessian.cpp,
accumulate_hessian_kernel.cu, gptq_cpp.cpp, gptq_cpp_kernel.cu) against the reference's vendored CUTLASS headers
(package/cutlass/include); nothing is copied into this repository and the only outputs are oracle/_ref/gptq_c.so and
ninja's object files beside it (oracle/_ref/ is git-ignored but travels to the GPU box).  The reference's setup.py
passes no arch flags (package/setup.py:24-30 are commented out), so the arch is whatever torch derives from
TORCH_CUDA_ARCH_LIST=10.0 (it emits compute_100a/sm_100a); the reference has no Blackwell-specific code, its CUTLASS
kernel is an sm80 mma.sync GEMM recompiled for the B200.

The built module is the GPU-side checker for a2 (CUTLASS Hessian kernel) and a6 (gptq_quantize_range):
tests/test_gpu_vs_reference_ext.py runs it beside libgptq_b200.so on identical inputs.

    python oracle/build_ref.py        # no-op (exit 0) when /root/reference is absent
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_PKG = '/root/reference/package'
OUT = os.path.join(HERE, '_ref')


def build(verbose: bool = False) -> str | None:
    src = os.path.join(REF_PKG, 'gptq')
    if not os.path.isdir(src):
        return None
    so = os.path.join(OUT, 'gptq_c.so')
    names = ['pybind.cpp', 'accumulate_hessian.cpp', 'accumulate_hessian_kernel.cu', 'gptq_cpp.cpp', 'gptq_cpp_kernel.cu']
    sources = [os.path.join(src, n) for n in names]
    if os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in sources):
        return so
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault('TORCH_CUDA_ARCH_LIST', '10.0')
    os.environ.setdefault('MAX_JOBS', str(os.cpu_count() or 4))
    from torch.utils import cpp_extension
    cpp_extension.load(
        name='gptq_c', sources=sources,
        extra_include_paths=[os.path.join(REF_PKG, 'cutlass', 'include'),
                             os.path.join(REF_PKG, 'cutlass', 'tools', 'util', 'include')],
        build_directory=OUT, verbose=verbose, is_python_module=False)
    return so if os.path.exists(so) else None


if __name__ == '__main__':
    p = build(verbose='-v' in sys.argv)
    print('oracle/_ref:', p or 'reference sources not present, nothing built')
