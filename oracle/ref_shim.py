"""Import the UNMODIFIED reference (ba, calib_ransac, util) in the build container.

Test infrastructure only.  /root/reference does not exist on the GPU box, so nothing
run there may import this; it is used by oracle/gen_golden.py (which writes
tests/golden/*.npz) and by tests that skip when the reference is absent.
Stubs follow SURVEY.md §8c: matplotlib / nvgpu / pycalib are not installed, and
np.int / np.alltrue are gone from NumPy 2 (np.bool is NOT aliased — it exists).
"""
import os
import sys
import types

import numpy as np

REF = os.environ.get("CALIB_REFERENCE_DIR", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF, "ba.py"))


def load():
    """Returns (ba, calib_ransac, util) modules of the reference."""
    if "ref_ba" in sys.modules:
        return sys.modules["ref_ba"], sys.modules["ref_calib_ransac"], sys.modules["ref_util"]
    for name, val in (("int", int), ("alltrue", np.all)):
        if not hasattr(np, name):
            setattr(np, name, val)

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    saved = {k: sys.modules.get(k) for k in ("util", "ba", "calib_ransac", "argument")}
    for k in saved:
        sys.modules.pop(k, None)
    stub("matplotlib"); stub("matplotlib.pyplot"); stub("matplotlib.pylab")
    stub("matplotlib.style", available=[])
    stub("mpl_toolkits"); stub("mpl_toolkits.mplot3d", Axes3D=object); stub("nvgpu")
    pc = stub("pycalib"); pc.calib = stub("pycalib.calib", triangulate=lambda *a: None)
    sys.path.insert(0, REF)
    argv = sys.argv
    sys.argv = ["x"]
    try:
        import util as ref_util
        sys.modules["pycalib.calib"].triangulate = ref_util.triangulate  # vendored copy, util.py:240-258
        import ba as ref_ba
        import calib_ransac as ref_cr
    finally:
        sys.argv = argv
        sys.path.remove(REF)
    # un-shadow: the repo has its own top-level ba.py / calib_ransac.py / argument.py
    for k in ("util", "ba", "calib_ransac", "argument"):
        sys.modules.pop(k, None)
        if saved[k] is not None:
            sys.modules[k] = saved[k]
    for k in ("matplotlib", "matplotlib.pyplot", "matplotlib.pylab", "matplotlib.style",
              "mpl_toolkits", "mpl_toolkits.mplot3d", "nvgpu", "pycalib", "pycalib.calib"):
        sys.modules.pop(k, None)
    sys.modules["ref_ba"], sys.modules["ref_calib_ransac"], sys.modules["ref_util"] = ref_ba, ref_cr, ref_util
    return ref_ba, ref_cr, ref_util
