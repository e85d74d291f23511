"""Importable alias of the package directory (whose contract-mandated name contains hyphens)."""
import importlib
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
_pkg = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200")
sys.modules[__name__] = _pkg
