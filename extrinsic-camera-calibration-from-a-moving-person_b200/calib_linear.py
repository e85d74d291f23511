"""Non-RANSAC linear calibration: drop-in for the reference's calib_linear.py.

The model fit is the same `calib_linear` as the RANSAC path's final inlier fit, run on ALL orientations and ray
sets (rsx_calib_linear with NULL masks: arrow-structured inverse subspace iteration, the 3(M+C) eigenproblem is
never densified as calib_linear.py:96-98 does), followed by the DLT re-triangulation of every point under the
fitted cameras (calib_linear.py:127-133, rsx_triangulate).  No CPU fallback.
"""
from __future__ import annotations

import os

import numpy as np

from . import dataio
from .ba import OP_BONE
from .calib_ransac import RansacProblem, ransac_inputs


def calib_linear(v_CxNx3, n_CxMx3):
    """reference calib_linear.calib_linear (calib_linear.py:48-135) -> R (C,3,3), t (C,3,1), X2 (M,3): unlike
    calib_ransac.calib_linear the returned points are re-triangulated, not the null-space block."""
    prob = RansacProblem(v_CxNx3, n_CxMx3)
    try:
        R, t, _, info = prob.calib_linear(want_X=False)
        w = np.array(info["eigs"])
        if w[3] / w[4] > 1e-4:                                                   # calib_linear.py:98-100
            print(f"WARN: degenerate case (only 4 eigenvalues should be zero): lambda={w}")
        X4 = prob.triangulate(R, t)
    finally:
        prob.close()
    return R, t, X4[:, :3].copy()


def main_linear(dirname, gid, aid, pid, bone_idx, joint_idx, bObs_mask, *, frame_skip=15):
    """reference calib_linear.main_linear (calib_linear.py:138-215).  `joint_idx` is accepted and ignored, as the
    reference's joints2projections ignores it (util.py:556-563).  -> (R, t, X, e, CAMID, K); `e` is, as in the
    reference (:207-213, the list E is never used), the mean reprojection error of the LAST camera only."""
    sub = "2d_joint_mask" if bObs_mask else "2d_joint"
    camid, K, _, _, _, p3d, s3d, p2d, s2d, _ = dataio.load_scene(dirname, gid, aid, pid, joint2d_dir=sub)
    k = frame_skip
    v, n = ransac_inputs(p3d[:, ::k], s3d[:, ::k], p2d[:, ::k], s2d[:, ::k], K, bone_idx)
    assert v.shape[0] == len(camid) and v.shape[2] == 3
    print(f"vc={v.shape}, n={n.shape}")
    R, t, X = calib_linear(v, n)
    y_last = (n[-1] @ K[-1].T)[:, :2]
    p = K[-1] @ (R[-1] @ X.T + t[-1])
    e = float(np.mean(np.linalg.norm(y_last - (p[:2] / p[2]).T, axis=1)))
    return R, t, X, e, camid, K


def linear_cli(argv=None):
    """`python calib_linear.py ...` (calib_linear.py:218-266): writes PREFIX/results/linear_<a>_<b>[_mask].json."""
    from .cli import parse_args
    a = parse_args(argv)
    prefix = a.prefix + "/" + a.target
    print(f"dataset={a.dataset}")
    os.makedirs(a.prefix + "/results/", exist_ok=True)
    parts = a.target.split("_")
    json_out = a.prefix + "/results/linear_" + parts[1] + "_" + parts[2] + ("_mask" if a.obs_mask else "") + ".json"
    R, t, _, _, camid, K = main_linear(prefix, a.gid, a.aid, a.pid, OP_BONE, None, bool(a.obs_mask),
                                       frame_skip=a.frame_skip)
    dataio.save_cameras(json_out, camid, K, R, t)
    print(" ")
    return json_out
