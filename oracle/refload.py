"""Import the UNMODIFIED reference modules.  TEST INFRASTRUCTURE ONLY.

Two places can hold them: ``/root/reference`` (authoring container only) and ``oracle/_ref/`` (a byte
copy staged by ``oracle/stage_ref.py``; git-ignored, travels to the GPU box with the snapshot, verified
against its ``MANIFEST.json`` before import).  Users: ``tests/golden/make_golden.py``, the oracle
self-checks in ``tests/`` and the timed CPU baseline of ``bench.py`` (``kind: "reference"``).
Recipe: SURVEY.md Appendix B -- ``model.py:12,14-15`` and ``libs/utils.py:8-9`` import ``matplotlib`` and
``gym_microrts`` at module top; empty stub modules let ``model`` and ``libs.utils`` import unchanged.
``parser`` / ``microbeast`` must not be imported (argparse at import time, JVM-backed env).
"""
import hashlib
import json
import os
import sys
import types

REFERENCE_DIR = "/root/reference"
STAGED_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _staged_ok() -> bool:
    try:
        with open(os.path.join(STAGED_DIR, "MANIFEST.json")) as f:
            want = json.load(f)["sha256"]
        for rel, digest in want.items():
            with open(os.path.join(STAGED_DIR, rel), "rb") as f:
                if hashlib.sha256(f.read()).hexdigest() != digest:
                    return False
        return bool(want)
    except Exception:
        return False


def location():
    """Directory the reference would be imported from, or ``None``."""
    if os.path.isfile(os.path.join(REFERENCE_DIR, "model.py")):
        return REFERENCE_DIR
    if _staged_ok():
        return STAGED_DIR
    return None


def available() -> bool:
    return location() is not None


def load():
    """Returns ``(model_module, utils_module)`` of the reference."""
    where = location()
    if where is None:
        raise RuntimeError("reference not present at " + REFERENCE_DIR + " nor staged in " + STAGED_DIR)
    for name in ("matplotlib", "matplotlib.pyplot", "gym_microrts",
                 "gym_microrts.envs", "gym_microrts.envs.vec_env"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["gym_microrts"].microrts_ai = types.SimpleNamespace()
    sys.modules["gym_microrts.envs.vec_env"].MicroRTSGridModeVecEnv = object
    if where not in sys.path:
        sys.path.insert(0, where)
    import importlib
    ref_model = importlib.import_module("model")
    ref_utils = importlib.import_module("libs.utils")
    return ref_model, ref_utils
