"""Stage the UNMODIFIED reference learner modules into ``oracle/_ref/``.  TEST INFRASTRUCTURE ONLY.

    python oracle/stage_ref.py            # run in the authoring container (needs /root/reference)

The reference is pure Python, so "building" it means placing the two modules the learner hot path
lives in -- ``model.py`` (``Agent``, ``CategoricalMasked``) and ``libs/utils.py`` (``PPO_learn``) --
where the GPU box can import them: ``/root/reference`` does not exist there, ``oracle/_ref/`` (git-ignored
like a built ``.so``, not gpurun-ignored) travels with the snapshot.  The files are byte copies; a
``MANIFEST.json`` with their sha256 is written beside them and ``oracle/refload.py`` refuses a staged copy
whose hash does not match it.  Nothing here is committed to history and nothing under ``microbeast_b200/``
reads it: its only users are ``bench.py --impl reference`` / the ``cpu_baseline`` leg (the timed CPU
baseline, ``kind: "reference"``) and the oracle self-checks in ``tests/``.
"""
import hashlib
import json
import os
import shutil
import sys

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
FILES = ("model.py", os.path.join("libs", "utils.py"))


def sha256(path: str) -> str:
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def stage(src: str = SRC, dst: str = DST) -> bool:
    """Copies the reference modules; returns False (and touches nothing) when ``src`` is absent."""
    if not all(os.path.isfile(os.path.join(src, f)) for f in FILES):
        return False
    manifest = {}
    for rel in FILES:
        out = os.path.join(dst, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), out)
        os.chmod(out, 0o644)
        manifest[rel] = sha256(out)
    with open(os.path.join(dst, "MANIFEST.json"), "w") as f:
        json.dump({"source": src, "sha256": manifest}, f, indent=1)
    return True


if __name__ == "__main__":
    ok = stage()
    print("staged reference into", DST if ok else "(nothing: /root/reference absent)")
    sys.exit(0)
