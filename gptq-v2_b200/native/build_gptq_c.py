"""Build gptq-v2_b200/lib/gptq_c.so: the reference's native module `gptq_c` (package/gptq/pybind.cpp) as a thin torch
extension over libgptq_b200.so.  Host C++ only (no device code): it forwards raw pointers through the C ABI.

    python gptq-v2_b200/native/build_gptq_c.py
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(os.path.dirname(HERE), 'lib')


def build(verbose: bool = False) -> str:
    src = os.path.join(HERE, 'gptq_c.cpp')
    out = os.path.join(LIB, 'gptq_c.so')
    core = os.path.join(LIB, 'libgptq_b200.so')
    hdr = os.path.join(os.path.dirname(os.path.dirname(HERE)), 'include', 'gptq_b200.h')
    if os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(f) for f in (src, hdr)):
        return out
    assert os.path.exists(core), 'build libgptq_b200.so first (make -C gptq-v2_b200/csrc)'
    os.environ.setdefault('TORCH_CUDA_ARCH_LIST', '10.0')
    os.environ.setdefault('MAX_JOBS', str(os.cpu_count() or 4))
    from torch.utils import cpp_extension
    bdir = os.path.join(os.path.dirname(HERE), 'build', 'gptq_c')
    os.makedirs(bdir, exist_ok=True)
    built = os.path.join(bdir, 'gptq_c.so')
    try:
        cpp_extension.load(name='gptq_c', sources=[src], build_directory=bdir, verbose=verbose, is_python_module=False, with_cuda=True,
                           extra_cflags=['-fvisibility=hidden'],
                           extra_ldflags=[f'-L{LIB}', '-lgptq_b200', "-Wl,-rpath,'$$ORIGIN'"])
    except OSError:
        # torch loads the fresh library from the build directory, where libgptq_b200.so is not beside it ($ORIGIN) and, in
        # the build container, no CUDA driver exists: the link step has succeeded if the file is there
        if not os.path.exists(built):
            raise
    shutil.copy2(built, out)
    return out


if __name__ == '__main__':
    print(build(verbose='-v' in sys.argv))
