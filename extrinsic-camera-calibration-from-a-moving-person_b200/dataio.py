"""JSON wire formats either side of the hot path, format-compatible with the reference's
loaders / savers (util.py:357-487, 668-678, 726-767; ba.py:213-232,333-340; calib_ransac.py:325-332):

  cameras_G%03d.json          {"CAMID": [C], "K": [C][3][3], "R_w2c": [C][3][3], "t_w2c": [C][3][1]}
  skeleton_w_G%03d.json       {"skeleton": [N][25][3], "frame_indices": [N]}
  {2d,3d}_joint/A%03d_P%03d_G%03d_C%03d.json
                              {"data": [{"frame_index": i, "skeleton": [{"pose": [25*2 | 25*3], "score": [25]}]}]}
  results/<target>[_ba].json  same keys as cameras_*.json (no dist)

A NumPy side channel (`scene.npz` next to cameras_*.json) is accepted as well: at BASELINE cfg4 sizes parsing
1e8 observations as JSON text would dwarf the GPU time.  Plain file IO, no GPU work.
"""
from __future__ import annotations

import functools
import json
import os

import numpy as np

N_SLOTS = 25


def joint_file(aid, pid, gid, cid):
    return f"A{aid:03d}_P{pid:03d}_G{gid:03d}_C{cid:03d}.json"


def load_cameras(path):
    """util.load_eldersim_camera (util.py:377-389)."""
    with open(path) as fp:
        d = json.load(fp)
    camid = np.asarray(d["CAMID"], dtype=np.int64)
    K = np.asarray(d["K"], dtype=np.float64)
    R = np.asarray(d["R_w2c"], dtype=np.float64)
    t = np.asarray(d["t_w2c"], dtype=np.float64)
    assert len(camid) == len(K) == len(R) == len(t)
    return camid, K, R, t


def save_cameras(path, camid, K, R_w2c, t_w2c):
    """The result writer shared by ba.py:333-340 and calib_ransac.py:325-332."""
    out = {"CAMID": np.asarray(camid).tolist(), "K": np.asarray(K).tolist(),
           "R_w2c": np.asarray(R_w2c).tolist(), "t_w2c": np.asarray(t_w2c).reshape(-1, 3, 1).tolist()}
    with open(path, "w") as fp:
        json.dump(out, fp, indent=2, ensure_ascii=True)


def load_skeleton_w(path):
    """util.load_eldersim_skeleton_w (util.py:392-400)."""
    with open(path) as fp:
        d = json.load(fp)
    p = np.asarray(d["skeleton"], dtype=np.float64)
    f = np.asarray(d["frame_indices"], dtype=np.int64)
    assert len(p) == len(f)
    return p, f


def save_skeleton_w(path, X_w):
    X_w = np.asarray(X_w, dtype=np.float64)
    with open(path, "w") as fp:
        json.dump({"skeleton": X_w.tolist(), "frame_indices": list(range(1, len(X_w) + 1))}, fp, indent=2, ensure_ascii=True)


def load_poses(path):
    """util.load_poses (util.py:357-374): first skeleton of every frame record."""
    with open(path) as fp:
        d = json.load(fp)
    rec = d["data"]
    frames = np.asarray([int(r["frame_index"]) for r in rec])
    pose = np.asarray([r["skeleton"][0]["pose"] for r in rec], dtype=np.float64)
    score = np.asarray([r["skeleton"][0]["score"] for r in rec], dtype=np.float64)
    return frames, pose, score


def save_poses(path, frames, pose, score):
    """ba.save_json (ba.py:213-232) / util.save_joint (util.py:726-752): NaN scores are written as JSON NaN
    exactly as Python's json module does for the reference."""
    data = [{"frame_index": int(f), "skeleton": [{"pose": np.asarray(p, dtype=np.float64).reshape(-1).tolist(),
                                                  "score": np.asarray(s, dtype=np.float64).tolist()}]}
            for f, p, s in zip(frames, pose, score)]
    with open(path, "w") as fp:
        json.dump({"data": data}, fp, indent=2, ensure_ascii=True)


def load_scene(dirname, gid, aid, pid, joint2d_dir="2d_joint"):
    """util.load_eldersim (util.py:420-487): cameras, world skeleton, per-camera 2-D / 3-D joints restricted to
    the frames present everywhere.  -> (CAMID, K, R_w2c, t_w2c, p3d_w, p3d, s3d, p2d, s2d, frames)"""
    npz = os.path.join(dirname, "scene.npz")
    camid, K, R, t = load_cameras(os.path.join(dirname, f"cameras_G{gid:03d}.json"))
    if os.path.isfile(npz):
        z = np.load(npz)
        return camid, K, R, t, z["p3d_w"], z["p3d"], z["s3d"], z["p2d"], z["s2d"], z["frames"]
    p3d_w, frames = load_skeleton_w(os.path.join(dirname, f"skeleton_w_G{gid:03d}.json"))
    f_all, p2, s2, p3, s3 = [], [], [], [], []
    for cid in camid:
        name = joint_file(aid, pid, gid, int(cid))
        f2d, a, b = load_poses(os.path.join(dirname, joint2d_dir, name))
        f3d, c, d = load_poses(os.path.join(dirname, "3d_joint", name))
        assert np.all(f2d == f3d)
        f_all.append(f2d); p2.append(a.reshape(-1, N_SLOTS, 2)); s2.append(b)
        p3.append(c.reshape(-1, N_SLOTS, 3)); s3.append(d)
    common = functools.reduce(np.intersect1d, [frames, *f_all])
    sel = [np.intersect1d(common, f, return_indices=True)[2] for f in f_all]
    pick = lambda arrs: np.asarray([a[i] for a, i in zip(arrs, sel)])
    return camid, K, R, t, p3d_w, pick(p3), pick(s3), pick(p2), pick(s2), common


def write_scene(dirname, sc, aid, pid, gid, *, npz=False):
    """Write a synthetic `scene.Scene` in the reference's directory layout (noise_<a>_<b>/ under a prefix)."""
    os.makedirs(os.path.join(dirname, "2d_joint"), exist_ok=True)
    os.makedirs(os.path.join(dirname, "3d_joint"), exist_ok=True)
    save_cameras(os.path.join(dirname, f"cameras_G{gid:03d}.json"), sc.camid, sc.K, sc.R_gt, sc.t_gt)
    frames = np.arange(1, sc.N + 1)
    if npz:
        np.savez(os.path.join(dirname, "scene.npz"), p3d_w=sc.X_w, p3d=sc.p3d, s3d=sc.s3d, p2d=sc.p2d, s2d=sc.s2d,
                 frames=frames)
        return
    save_skeleton_w(os.path.join(dirname, f"skeleton_w_G{gid:03d}.json"), sc.X_w)
    for i, cid in enumerate(sc.camid):
        name = joint_file(aid, pid, gid, int(cid))
        save_poses(os.path.join(dirname, "2d_joint", name), frames, sc.p2d[i], sc.s2d[i])
        save_poses(os.path.join(dirname, "3d_joint", name), frames, sc.p3d[i], sc.s3d[i])
