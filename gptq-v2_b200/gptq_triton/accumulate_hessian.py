"""Import-compatible stand-in for the reference's Triton module
(gptq_triton/accumulate_hessian.py:134-149, imported by gptq_py.py:6).  No Triton here:
the call goes to the tcgen05 SYRK in libgptq_b200.so."""
import torch

from gptq import hessian_syrk


def accumulate_hessian(mat_hessian: torch.Tensor, mat_input: torch.Tensor, save_lower_only: bool,
                       compute_lower_only: bool = True) -> None:
    """H += X^T X in place.  The reference computes lower+diagonal tiles and, unless
    save_lower_only, mirrors them; this build computes the upper triangle and always
    mirrors, so H is fully symmetric on return for every flag combination (a superset of
    what save_lower_only=True guarantees)."""
    assert compute_lower_only or not save_lower_only, \
        'compute_lower_only must be True when save_lower_only is True'
    hessian_syrk(mat_hessian, mat_input, mode=1)
