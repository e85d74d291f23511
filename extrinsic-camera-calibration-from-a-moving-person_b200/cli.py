"""Command-line surface of the hot path: the flags, file names and printed lines of the reference's
`ba.py` / `calib_ransac.py` __main__ blocks (ba.py:269-341, calib_ransac.py:280-334, argument.py:7-58), so that
`ba.sh` / `calib_ransac.sh` (and run_all.sh around them) keep working when the two scripts are swapped for the
ones at this repo's root.  cwd must hold config/config.yaml, as for the reference.
"""
from __future__ import annotations

import argparse
import os

import numpy as np
import yaml

from . import ba as _ba
from . import calib_ransac as _cr
from . import dataio


def _truthy(text):
    """distutils.util.strtobool semantics (argument.py:4,43-44) without distutils (gone from Python 3.12)."""
    t = str(text).strip().lower()
    if t in ("y", "yes", "t", "true", "on", "1"):
        return 1
    if t in ("n", "no", "f", "false", "off", "0"):
        return 0
    raise argparse.ArgumentTypeError(f"invalid truth value {text!r}")


def parse_args(argv=None):
    """Every flag of the reference's shared parser (argument.py:7-58) is accepted with its default; the hot path
    reads --prefix --aid --pid --gid --frame_skip --ba_lambda1 --ba_lambda2 --target --dataset --obs_mask
    --th_obs_mask --save_obs_mask --ransac_niter --ransac_seed."""
    p = argparse.ArgumentParser(description="human_calib (B200 hot path)")
    for name, typ, default in (
            ("dataset", str, "SynADL"), ("aid", int, 23), ("pid", int, 102), ("gid", int, 3), ("cid", int, 21),
            ("frame_skip", int, 15), ("noise_level", int, 0), ("prefix", str, "./third_party/SynADL/"),
            ("ransac_th_2d", float, 50.0), ("ransac_th_3d", float, 5.0), ("ransac_niter", int, 100),
            ("ransac_seed", int, 0), ("target", str, "noise_3_0"), ("th_truncate", float, 10.0),
            ("ba_lambda1", float, 1.0), ("ba_lambda2", float, 1.0), ("trials", int, 15),
            ("model", str, "pretrained_h36m_detectron_coco.bin"), ("retrain_pose", str, "linear_77_0_ba"),
            ("retrain_resume", str, "pretrained_h36m_detectron_coco.bin"), ("retrain_target_epoch", int, 100),
            ("obs_mask", _truthy, False), ("save_obs_mask", _truthy, False), ("th_obs_mask", int, 20),
            ("src_original", str, "./third_party/GAFA/lab/1013_2"), ("vis_type", str, "2d")):
        p.add_argument("--" + name, type=typ, default=default)
    p.add_argument("--calib", nargs="+", type=int, default=[0])
    return p.parse_args(argv)


def load_config(dataset, path="./config/config.yaml"):
    with open(path) as fp:
        return yaml.safe_load(fp.read())[dataset]


# ----------------------------------------------------------------------------- ba.py
def project_points(R, t, K, X, width, height):
    """util.project_cv2 (util.py:798-818): pinhole projection without distortion; projections left of / above the
    image or beyond width / height become NaN.  X (N, J, 3) -> (C, N, J, 2)."""
    out = []
    for Rc, tc, Kc in zip(R, t, K):
        q = (X.reshape(-1, 3) @ Rc.T + tc.reshape(1, 3)) @ Kc.T
        x = q[:, :2] / q[:, 2:3]
        bad = np.any(x < 0, axis=1) | (x[:, 0] > width) | (x[:, 1] > height)
        x[bad] = np.nan
        out.append(x.reshape(X.shape[0], X.shape[1], 2))
    return np.asarray(out)


def save_mask(prefix, gid, aid, pid, K, R, t, obs_mask, th_mask, width, height):
    """ba.save_mask (ba.py:235-266): re-load ALL frames and joints, triangulate with the optimised cameras (GPU),
    re-project, set the score of detections further than th_mask px from their re-projection to NaN and write
    2d_joint_mask[_ba]/."""
    print("save obs. mask")
    camid, _, _, _, _, _, _, p2d, s2d, frames = dataio.load_scene(prefix, gid, aid, pid)
    prob = _ba.BAProblem(K, p2d, s2d, None, None, _ba.OP_BONE, 0.0, 0.0)
    try:
        x_all = prob.triangulate(R, t)
    finally:
        prob.close()
    proj = project_points(R, t, K, x_all, width, height)
    with np.errstate(invalid="ignore"):
        far = np.linalg.norm(p2d - proj, axis=3) > th_mask
    s2d = np.array(s2d, dtype=np.float64)
    s2d[far] = np.nan
    sub = "2d_joint_mask_ba" if obs_mask else "2d_joint_mask"
    os.makedirs(os.path.join(prefix, sub), exist_ok=True)
    for i, cid in enumerate(camid):
        dataio.save_poses(os.path.join(prefix, sub, dataio.joint_file(aid, pid, gid, int(cid))), frames, p2d[i], s2d[i])
    return int(far.sum())


def ba_cli(argv=None):
    """`python ba.py ...` (ba.py:269-341)."""
    a = parse_args(argv)
    parts = a.target.split("_")
    prefix = a.prefix + "/" + "noise_" + parts[1] + "_" + parts[2]
    tag = a.target + ("_mask" if a.obs_mask else "")
    json_in = a.prefix + "/results/" + tag + ".json"
    json_out = a.prefix + "/results/" + tag + "_ba.json"
    cfg = load_config(a.dataset)
    joints = cfg["available_joints"]
    # NB: as in the reference (ba.py:304-310), --obs_mask only renames JSON_IN / JSON_OUT; 2d_joint/ is read either way
    s_camid, _, _, _, p3d_w, p3d, s3d, p2d, s2d, frames = dataio.load_scene(prefix, a.gid, a.aid, a.pid)
    camid, K, R, t = dataio.load_cameras(json_in)
    assert np.all(camid == s_camid)
    k = a.frame_skip
    p3d, s3d = p3d[:, ::k], s3d[:, ::k]
    p2d, s2d = p2d[:, ::k][:, :, joints, :], s2d[:, ::k][:, :, joints]
    print(f"dataset={a.dataset}")
    print(f"target BA={json_in}")
    R_opt, t_opt, _ = _ba.ba_main(camid, K, R, t, np.ascontiguousarray(p2d), np.ascontiguousarray(s2d),
                                  np.ascontiguousarray(p3d), np.ascontiguousarray(s3d), a.ba_lambda1, a.ba_lambda2)
    if a.save_obs_mask:
        save_mask(prefix, a.gid, a.aid, a.pid, K, R_opt, t_opt, a.obs_mask, a.th_obs_mask, cfg["width"], cfg["height"])
    dataio.save_cameras(json_out, camid, K, R_opt, t_opt)
    print(" ")
    return json_out


# ----------------------------------------------------------------------------- calib_ransac.py
def main_ransac(dirname, gid, aid, pid, bone_idx, joint_idx, *, frame_skip=15, n_iter=10, th_2d=3, th_3d=np.pi / 8,
                seed, group=None):
    """Drop-in for reference calib_ransac.main_ransac (calib_ransac.py:212-277); `joint_idx` is accepted and
    ignored, as the reference's joints2projections ignores it (util.py:556-563)."""
    camid, K, _, _, _, p3d, s3d, p2d, s2d, _ = dataio.load_scene(dirname, gid, aid, pid)
    k = frame_skip
    v, n = _cr.ransac_inputs(p3d[:, ::k], s3d[:, ::k], p2d[:, ::k], s2d[:, ::k], K, bone_idx)
    assert v.shape[0] == len(camid) and v.shape[2] == 3 and n.shape[2] == 3
    print(f"vc={v.shape}, n={n.shape}")
    R, t, _ = _cr.calib_linear_ransac(v, n, K, n_iter, th_2d, th_3d, seed, group=group)
    return R, t, None, None, camid, K


def ransac_cli(argv=None):
    """`python calib_ransac.py ...` (calib_ransac.py:280-334).  Thresholds come from config.yaml, not from the
    --ransac_th_* flags, exactly as in the reference (:305-309)."""
    a = parse_args(argv)
    prefix = a.prefix + "/" + a.target
    os.makedirs(a.prefix + "/results/", exist_ok=True)
    parts = a.target.split("_")
    json_out = a.prefix + "/results/ransac_" + parts[1] + "_" + parts[2] + ".json"
    print(f"dataset={a.dataset}")
    cfg = load_config(a.dataset)
    th_2d = cfg["ransac_th_2d"]
    th_3d = cfg["ransac_th_3d"] / 180 * np.pi
    R, t, _, _, camid, K = main_ransac(prefix, a.gid, a.aid, a.pid, _ba.OP_BONE, None, frame_skip=a.frame_skip,
                                       n_iter=a.ransac_niter, th_2d=th_2d, th_3d=th_3d, seed=a.ransac_seed)
    dataio.save_cameras(json_out, camid, K, R, t)
    print(" ")
    return json_out
