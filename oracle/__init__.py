"""CPU oracle for the microbeast learner hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on CPU torch/numpy, the algorithm of the reference
(Neos-codes/microbeast) for the one path this repository accelerates:

* the GridNet masked multi-discrete policy head
  (reference ``model.py:33-52`` ``CategoricalMasked`` and ``model.py:165-216``
  ``Agent.get_action``), and
* the V-trace targets / advantages and the pg / baseline / entropy loss
  reduction (reference ``libs/utils.py:277-329`` inside ``PPO_learn``).

Nothing under ``microbeast_b200/`` may import this package.  The only legal
importers are ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline``
/ ``--impl reference`` legs of ``bench.py``, and there only as the checker or
as the timed CPU baseline -- never as the product path.

Third-party arithmetic: the reference's distribution math lives in
``torch.distributions.Categorical`` (un-pinned by the reference, which ships no
requirements file).  The de-facto pin is torch 2.11.0, the version in this
image and on the GPU box.  ``oracle.head.MaskedCategorical`` subclasses that
very class, so the oracle executes the same third-party code as the reference.

Parity pinning: the reference has no golden vectors, known-answer tests or
fixtures for this path (its ``tests/`` are print-only scripts).  The oracle is
pinned instead against outputs of the *unmodified* reference run in the
authoring container: ``tests/golden/make_golden.py`` imports
``/root/reference`` (stub ``gym_microrts`` / ``matplotlib`` modules), runs the
real ``CategoricalMasked``, ``Agent.get_action`` and ``PPO_learn`` and commits
their outputs as small fixtures under ``tests/golden/``.  ``tests/test_oracle_*``
check every oracle function against those fixtures.

The inverse-CDF sampler is *defined* here (``oracle.head.sample_invcdf``): the
reference's sampler, ``Categorical.sample`` -> ``torch.multinomial`` on torch's
global RNG, has no hook for caller-supplied uniforms, so there is nothing in
the reference to be bit-exact with.  The definition is built on the reference
distribution class evaluated in fp64 and is cross-checked against the real
``.sample()`` at distribution level (chi-square) by the golden script.
"""

NVEC_CELL = (6, 4, 4, 4, 4, 7, 49)   # reference microbeast.py:145 / libs/utils.py:40
LOGITS_PER_CELL = 78
COMPONENTS_PER_CELL = 7
