"""microbeast_b200 -- B200-native (sm_100a) learner hot path of Neos-codes/microbeast.

Drop-in surface (same names / signatures as the reference):

* ``microbeast_b200.model.Agent``            <- reference ``model.py:112-220``
* ``microbeast_b200.model.CategoricalMasked`` (multi-discrete, whole frame)  <- ``model.py:33-52``
* ``microbeast_b200.learner.PPO_learn``      <- reference ``libs/utils.py:223-342``
* ``microbeast_b200.buffers.create_buffers`` / ``get_batch``  <- ``libs/utils.py:29-55,166-218``

The hot ops are hand-written CUDA kernels behind the C ABI of ``include/mbeast.h``
(``libmbeast_sm100.so``); importing the package loads that library and fails loudly
when it is missing.  PyTorch supplies device memory, streams, the conv/linear encoder
and ``torch.distributed``.
"""
from . import _lib                      # noqa: F401  (raises if the CUDA extension is missing)
from ._lib import ABI_VERSION, set_tuning

NVEC_CELL = (6, 4, 4, 4, 4, 7, 49)
LOGITS_PER_CELL = 78
COMPONENTS_PER_CELL = 7

__all__ = ["ABI_VERSION", "set_tuning", "NVEC_CELL", "LOGITS_PER_CELL", "COMPONENTS_PER_CELL"]
