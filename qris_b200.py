"""Importable alias for the package directory
``quantum-distributed-machine-learning-ris-channel-estimation_b200`` (its name, fixed by the project
layout, contains hyphens and so cannot appear in an ``import`` statement).  ``import qris_b200`` and
``from qris_b200 import estimators`` work from the repository root."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("quantum-distributed-machine-learning-ris-channel-estimation_b200")
sys.modules[__name__] = _pkg
