"""Loader for oracle/_ref/gptq_c.so -- the reference's own CUDA extension built by oracle/build_ref.py
(test infrastructure; never imported by the product)."""
import importlib.util
import os

SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'oracle', '_ref', 'gptq_c.so')
_mod = None


def load():
    """Returns the pybind module `gptq_c` (pybind.cpp:31-35) or None when it was not built."""
    global _mod
    if _mod is not None:
        return _mod
    if not os.path.exists(SO):
        return None
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    # spec name 'reference_ext.gptq_c': the init symbol is still PyInit_gptq_c, but pybind11 (3.x) caches module objects per
    # interpreter by spec NAME, so under the plain name this module and the repository's own lib/gptq_c.so would alias
    spec = importlib.util.spec_from_file_location('reference_ext.gptq_c', SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _mod = mod
    return mod
