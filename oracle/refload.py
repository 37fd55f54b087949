"""Import the UNMODIFIED reference modules (authoring container only).  TEST INFRASTRUCTURE ONLY.

``/root/reference`` does not exist on the GPU box; only ``tests/golden/make_golden.py``
and the CPU tests that are skipped when the directory is absent use this.
Recipe: SURVEY.md Appendix B -- ``model.py:12,14-15`` and ``libs/utils.py:8-9``
import ``matplotlib`` and ``gym_microrts`` at module top; empty stub modules
let ``model`` and ``libs.utils`` import unchanged.  ``parser`` / ``microbeast``
must not be imported (argparse at import time, JVM-backed env).
"""
import os
import sys
import types

REFERENCE_DIR = "/root/reference"


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "model.py"))


def load():
    """Returns ``(model_module, utils_module)`` of the reference."""
    if not available():
        raise RuntimeError("reference checkout not present at " + REFERENCE_DIR)
    for name in ("matplotlib", "matplotlib.pyplot", "gym_microrts",
                 "gym_microrts.envs", "gym_microrts.envs.vec_env"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["gym_microrts"].microrts_ai = types.SimpleNamespace()
    sys.modules["gym_microrts.envs.vec_env"].MicroRTSGridModeVecEnv = object
    if REFERENCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_DIR)
    import importlib
    ref_model = importlib.import_module("model")
    ref_utils = importlib.import_module("libs.utils")
    return ref_model, ref_utils
