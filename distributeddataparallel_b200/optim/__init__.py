"""Optimizers: ``SGD`` (what /root/reference/dpp.py:41 uses) and ``Adam``.

Both subclass the torch optimizers for their *bookkeeping* -- param groups, ``state_dict``
format (SURVEY.md §5.4 lists the exact dict the reference's ``optim.SGD`` produces),
``zero_grad`` -- and replace ``step()`` on CUDA with ONE launch of a fused sm_100a
multi-tensor kernel (csrc/kernels/optim.cu) driven by a device-resident pointer table,
instead of the ``_foreach_*`` chains (K20).
"""
from .sgd import SGD
from .adam import Adam

__all__ = ["SGD", "Adam"]
