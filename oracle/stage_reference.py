"""Stage the reference's own ``Estimators_QuantumNAT_onchipQNN.py`` for the GPU tests.  TEST INFRASTRUCTURE ONLY.

The reference is pure Python, so its counterpart of "compile the reference's sources into ``oracle/_ref/``"
is this: where ``/root/reference`` exists (the build container), ``__graft_entry__.build()`` calls
``stage()`` and the file is placed, unmodified and byte for byte, in ``oracle/_ref/`` next to its SHA-256.
``oracle/_ref/`` is git-ignored (no reference source ever enters this repository's history) but travels to
the GPU box with the snapshot, where ``tests/test_module_gpu.py`` imports it with ``pennylane_shim``
installed as ``pennylane`` and runs the reference's own ``QSC_P128.forward``
(Estimators_QuantumNAT_onchipQNN.py:167-203) on the CUDA kernels.  Nothing in the product reads this
directory.
"""
import hashlib
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAME = "Estimators_QuantumNAT_onchipQNN.py"
SOURCE = os.path.join("/root/reference", NAME)
STAGED_DIR = os.path.join(ROOT, "oracle", "_ref")
STAGED = os.path.join(STAGED_DIR, NAME)


def sha256(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def stage() -> bool:
    """Copy the reference file into oracle/_ref/ if the reference checkout is present.  Returns whether
    a staged copy exists afterwards."""
    if os.path.exists(SOURCE):
        os.makedirs(STAGED_DIR, exist_ok=True)
        shutil.copyfile(SOURCE, STAGED)
        with open(STAGED + ".sha256", "w") as f:
            f.write(sha256(SOURCE) + "\n")
    return os.path.exists(STAGED)


def locate():
    """Path of the reference's estimator file: the checkout itself where it exists, else the staged copy
    (checked against the recorded digest), else None."""
    if os.path.exists(SOURCE):
        return SOURCE
    if os.path.exists(STAGED) and os.path.exists(STAGED + ".sha256"):
        if open(STAGED + ".sha256").read().strip() == sha256(STAGED):
            return STAGED
    return None


if __name__ == "__main__":
    print("staged" if stage() else "reference checkout not present; nothing staged")
