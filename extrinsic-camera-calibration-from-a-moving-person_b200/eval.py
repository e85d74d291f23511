"""Evaluation against ground truth: drop-in for the reference's eval.py.

`eval_main` keeps the reference's signature, printed table and JSON summary (eval.py:111-173).  The two passes over
all observations run on the GPU through the C ABI — the weighted triangulation of every (frame, joint)
(bax_triangulate == util.triangulate_with_conf, eval.py:134) and the reprojection error of every observation
(bax_reproject_error == eval.calc_repro_error, eval.py:76-108); the similarity alignment of C camera centres and the
per-camera E_R / E_t are O(C) host arithmetic, as in the reference.

Third-party arithmetic: `pycalib.calib.absolute_orientation` (pycalib-simple 2022.9.2.2, eval.py:9,45) is NOT under
/root/reference and not installable here.  `absolute_orientation` below restates its published algorithm (centre both
point sets, scale = ratio of the summed centred norms, rotation by the orthogonal Procrustes SVD with the determinant
fix, translation from the means).  **Parity of the alignment is unpinned**: no reference fixture or run exists for it;
the tests check its defining property (an exact similarity transform is recovered) instead.
"""
from __future__ import annotations

import json
import os

import numpy as np

from . import ba as _ba
from . import dataio


def absolute_orientation(p, q, no_scaling=False):
    """R, t, s with q ~= s R p + t for 3xN point sets (pycalib.calib.absolute_orientation, restated)."""
    p = np.asarray(p, dtype=np.float64); q = np.asarray(q, dtype=np.float64)
    assert p.ndim == q.ndim == 2 and p.shape[0] == q.shape[0] == 3 and p.shape[1] == q.shape[1]
    mp, mq = p.mean(axis=1), q.mean(axis=1)
    pc, qc = p - mp[:, None], q - mq[:, None]
    s = 1.0 if no_scaling else np.sum(np.linalg.norm(qc, axis=0)) / np.sum(np.linalg.norm(pc, axis=0))
    u, _, vt = np.linalg.svd(qc @ (s * pc).T)
    R = u @ np.diag([1.0, 1.0, np.linalg.det(u @ vt)]) @ vt
    t = mq - s * (R @ mp)
    return R, t, s


def eval_R(R1, R2):
    """util.eval_R (util.py:183-199): rotation angle of R1^T R2, divided by sqrt(2)."""
    e = (np.trace(R1.T @ R2) - 1.0) / 2.0
    return abs(np.arccos(min(1.0, max(-1.0, e)))) / np.sqrt(2)


def eval_t(t1, t2):
    """util.eval_t (util.py:202-203)."""
    return np.linalg.norm(np.asarray(t1).flatten() - np.asarray(t2).flatten())


def align(gR_w2c, gt_w2c, R_w2c, t_w2c, X3D):
    """eval.align (eval.py:29-63): similarity transform that maps the estimated camera centres onto the ground-truth
    ones, applied to the cameras and to the points.  -> (Rw2c, tw2c (C,3,1), R2n, t2n, x3d_new (n,3))."""
    assert X3D.shape[1] == 3
    assert np.all(gR_w2c.shape == R_w2c.shape) and np.all(gt_w2c.shape == t_w2c.shape)
    Cn = len(R_w2c)
    p1 = np.array([-gR_w2c[i].T @ gt_w2c[i] for i in range(Cn)]).reshape(-1, 3).T
    p2 = np.array([-R_w2c[i].T @ t_w2c[i] for i in range(Cn)]).reshape(-1, 3).T
    R, t, s = absolute_orientation(p2, p1)
    x3d_new = s * R @ X3D.T + t[:, None]
    R2n = np.array([R_w2c[i] @ R.T for i in range(Cn)])
    Rc2w = np.array([R @ R_w2c[i].T for i in range(Cn)])
    t2n = np.array([s * R @ (-R_w2c[i].T @ t_w2c[i].reshape(3, 1)) + t.reshape(3, 1) for i in range(Cn)])
    # util.invRT_batch (util.py:584-612): the inverse of the 4x4 [Rc2w | c; 0 1] by np.linalg.inv — NOT a transpose:
    # linear_*/ransac_* cameras carry non-orthonormal R (calib_ransac.py:64-70) and run_all.sh evaluates them too
    Rw2c, tw2c = np.empty((Cn, 3, 3)), np.empty((Cn, 3))
    for i in range(Cn):
        T = np.eye(4)
        T[:3, :3], T[:3, 3] = Rc2w[i], t2n[i].reshape(3)
        Ti = np.linalg.inv(T)
        Rw2c[i], tw2c[i] = Ti[:3, :3], Ti[:3, 3]
    return Rw2c, tw2c[:, :, None], R2n, t2n, x3d_new.T


def calc_repro_error(mask_2ds, frame_skip, x_all, gp2d, gK, height, width, R_w2c, t_w2c, gs2d=None):
    """eval.calc_repro_error (eval.py:76-108) -> (E_p, x_gt_all, x_est_all, None): per-observation pixel errors of
    every frame_skip-th frame where mask_2ds is set (GPU).  x_est_all is rebuilt on the host only for callers that
    plot it; the homogeneous list of the reference is not kept."""
    mask_2ds = np.asarray(mask_2ds, dtype=bool)
    Cn, F, J = mask_2ds.shape
    prob = _ba.BAProblem(gK, np.asarray(gp2d, dtype=np.float64), mask_2ds.astype(np.float64), None, None, _ba.OP_BONE, 0.0, 0.0)
    try:
        e_all = prob.reproject_error(R_w2c, t_w2c, np.asarray(x_all, dtype=np.float64).reshape(F, J, 3))
    finally:
        prob.close()
    m = mask_2ds[:, ::frame_skip]
    E_p = np.concatenate([e_all[i, ::frame_skip][m[i]] for i in range(Cn)])
    x_gt_all = np.full((Cn, F, J, 2), np.nan)
    for i in range(Cn):
        x_gt_all[i, ::frame_skip][m[i]] = np.asarray(gp2d)[i, ::frame_skip][m[i]]
    return E_p, x_gt_all, None, None


def create_val_summary(val):
    return {"min": f"{np.min(val):.04f}", "median": f"{np.median(val):.04f}", "mean": f"{np.mean(val):.04f}",
            "max": f"{np.max(val):.04f}"}


def eval_metrics(gK, R_w2c, t_w2c, gR_w2c, gt_w2c, gp2d, gs2d, gp3d_w, frame_skip, scale):
    """The four error vectors of eval_main (eval.py:126-150): E_R (C,), E_t (C,), E_p (visible observations of every
    frame_skip-th frame, camera-major), E_x (joints with ground truth)."""
    gp2d = np.asarray(gp2d, dtype=np.float64); gs2d = np.asarray(gs2d, dtype=np.float64)
    Cn, N, J = gs2d.shape
    prob = _ba.BAProblem(gK, gp2d, gs2d, None, None, _ba.OP_BONE, 0.0, 0.0)
    try:
        x_all = prob.triangulate(R_w2c, t_w2c)                          # (N, J, 3), GPU
        e_all = prob.reproject_error(R_w2c, t_w2c, x_all)               # (C, N, J), GPU
    finally:
        prob.close()
    nR, nt, _, _, np3d_w = align(gR_w2c, gt_w2c, R_w2c, t_w2c, x_all.reshape(-1, 3))
    np3d_w = np3d_w.reshape(N, J, 3)
    mask_3dw = np.all(~np.isnan(gp3d_w), axis=2)
    E_x = np.linalg.norm(np3d_w[mask_3dw] - gp3d_w[mask_3dw], axis=1) * scale
    E_R = np.array([eval_R(gR_w2c[i], nR[i]) for i in range(Cn)])
    E_t = np.array([eval_t(gt_w2c[i], nt[i]) for i in range(Cn)]) * scale
    mask = (gs2d > 0)[:, ::frame_skip]
    E_p = np.concatenate([e_all[i, ::frame_skip][mask[i]] for i in range(Cn)])
    return E_R, E_t, E_p, E_x


def eval_main(gK, R_w2c, t_w2c, gR_w2c, gt_w2c, gp2d, gs2d, gp3d_w, frame_skip, width, height, json_out, scale):
    """reference eval.eval_main (eval.py:111-173): same arguments, printed table and JSON file."""
    E_R, E_t, E_p, E_x = eval_metrics(gK, R_w2c, t_w2c, gR_w2c, gt_w2c, gp2d, gs2d, gp3d_w, frame_skip, scale)
    print("       Min  Median  Mean  Max")
    for name, v in (("E_R", E_R), ("E_t", E_t), ("E_p", E_p), ("E_x", E_x)):
        print(f"{name}: {np.min(v):.04f} {np.median(v):.04f} {np.mean(v):.04f} {np.max(v):.04f}")
    with open(json_out, "w") as f:
        json.dump({"E_R": create_val_summary(E_R), "E_t": create_val_summary(E_t), "E_p": create_val_summary(E_p),
                   "E_x": create_val_summary(E_x)}, f)
    return E_R, E_t, E_p, E_x


def eval_cli(argv=None):
    """`python eval.py ...` (eval.py:176-232): ground truth from PREFIX/gt_subset, cameras from
    PREFIX/results/<target>.json, summary to PREFIX/results/eval/<target>.json."""
    from .cli import load_config, parse_args
    a = parse_args(argv)
    gt_dir = a.prefix + "/gt_subset"
    json_in = a.prefix + "/results/" + a.target + ".json"
    os.makedirs(a.prefix + "/results/eval/", exist_ok=True)
    json_out = a.prefix + "/results/eval/" + a.target + ".json"
    print(f"target={json_in}")
    cfg = load_config(a.dataset)
    _, gK, gR, gt, gp3d_w, _, _, gp2d, gs2d, _ = dataio.load_scene(gt_dir, a.gid, a.aid, a.pid)
    _, _, R, t = dataio.load_cameras(json_in)
    eval_main(gK, R, t, gR, gt, gp2d, gs2d, gp3d_w, a.frame_skip, cfg["width"], cfg["height"], json_out, cfg["scale"])
    print(" ")
    return json_out
