"""Importable alias of the package directory (whose contract-mandated name contains hyphens).

`import calib_b200` binds the package and every submodule under both names to ONE module object each, so
`calib_b200.ba` and `extrinsic-camera-calibration-from-a-moving-person_b200.ba` can never diverge."""
import importlib
import os
import sys

_NAME = "extrinsic-camera-calibration-from-a-moving-person_b200"
_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
_pkg = importlib.import_module(_NAME)
_first = ("_lib", "_rsx_sigs", "scene", "ba", "calib_ransac", "dataio", "cli")
_rest = sorted(f[:-3] for f in os.listdir(os.path.join(_here, _NAME)) if f.endswith(".py") and f != "__init__.py")
for _sub in _first + tuple(s for s in _rest if s not in _first):       # every module of the package, one object each
    if os.path.isfile(os.path.join(_here, _NAME, _sub + ".py")):
        _m = importlib.import_module(_NAME + "." + _sub)
        sys.modules["calib_b200." + _sub] = _m
sys.modules["calib_b200"] = _pkg
