"""Seeded synthetic scenes of the shapes named in BASELINE.json.

The reference ships no data (its SynADL subset is a download link), so every
test, golden vector and bench line in this repo is driven by this generator.
It produces the arrays `ba.ba_main` / `calib_ransac.main_ransac` receive after
`util.load_eldersim` (reference util.py:420-487): 25 OpenPose joint slots per
frame, per-camera 2-D detections + scores and per-camera 3-D poses + scores.

Everything is float64 and NumPy-only so the oracle, the tests and the CUDA
path are fed bit-identical inputs.
"""
from __future__ import annotations

import dataclasses

import numpy as np

N_SLOTS = 25  # OpenPose-25 joint axis carried by every file (reference util.py:443,452)


@dataclasses.dataclass
class Scene:
    camid: np.ndarray   # (C,) int
    K: np.ndarray       # (C,3,3)
    R_gt: np.ndarray    # (C,3,3) world->camera
    t_gt: np.ndarray    # (C,3,1)
    R_init: np.ndarray  # (C,3,3) perturbed start for BA
    t_init: np.ndarray  # (C,3,1)
    X_w: np.ndarray     # (N,25,3) world joints
    p2d: np.ndarray     # (C,N,25,2)
    s2d: np.ndarray     # (C,N,25)
    p3d: np.ndarray     # (C,N,25,3) camera-frame joints (pose-network output stand-in)
    s3d: np.ndarray     # (C,N,25)
    seed: int
    width: int
    height: int

    @property
    def C(self):
        return self.K.shape[0]

    @property
    def N(self):
        return self.p2d.shape[1]


def rodrigues(rvec):
    """Axis-angle -> rotation matrix, textbook formula (matches cv2.Rodrigues to 1e-14)."""
    rvec = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = np.linalg.norm(rvec)
    if th < 1e-300:
        return np.eye(3)
    k = rvec / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * Kx


def ring_cameras(C, radius=4.0, height=-1.0):
    """C cameras on a ring looking at the origin; rows of R_w2c are the camera axes."""
    R = np.zeros((C, 3, 3))
    t = np.zeros((C, 3, 1))
    for i in range(C):
        a = 2.0 * np.pi * i / C
        c = np.array([radius * np.cos(a), height, radius * np.sin(a)])
        z = -c / np.linalg.norm(c)
        up = np.array([0.0, 1.0, 0.0])
        x = np.cross(up, z)
        x /= np.linalg.norm(x)
        y = np.cross(z, x)
        R[i] = np.stack([x, y, z])
        t[i, :, 0] = -R[i] @ c
    return R, t


def make_scene(C=4, N=1000, seed=0, *, sigma_px=3.0, sigma_3d=0.03, score_mode="ones",
               p_drop_2d=0.0, p_drop_3d=0.0, width=640, height=360, focal=320.0,
               init_rot_sigma=0.02, init_t_sigma=0.05, camid=None):
    """Build one scene.  `score_mode`: "ones" (SynADL-like) or "uniform" (detector-like U(0.2,1)).

    Partial visibility follows SURVEY.md §8d: a dropped 2-D detection gets score 0
    (never NaN) and every (frame, joint) keeps >= 2 views, otherwise the reference
    aborts in `triangulate_with_conf` (util.py:344-349).
    """
    rng = np.random.default_rng(seed)
    R_gt, t_gt = ring_cameras(C)
    K = np.tile(np.array([[focal, 0, width / 2.0], [0, focal, height / 2.0], [0, 0, 1.0]]), (C, 1, 1))

    # person: root random walk + fixed skeleton offsets + per-frame articulation
    root = np.cumsum(rng.normal(0, 0.05, (N, 1, 3)), axis=0)
    root -= root.mean(axis=0, keepdims=True)
    root = np.clip(root, -1.0, 1.0)
    skel = rng.normal(0, 0.3, (1, N_SLOTS, 3))
    X_w = root + skel + rng.normal(0, 0.05, (N, N_SLOTS, 3))

    p2d = np.zeros((C, N, N_SLOTS, 2))
    p3d = np.zeros((C, N, N_SLOTS, 3))
    for c in range(C):
        Xc = X_w @ R_gt[c].T + t_gt[c, :, 0]
        q = Xc @ K[c].T
        p2d[c] = q[..., :2] / q[..., 2:3] + rng.normal(0, sigma_px, (N, N_SLOTS, 2))
        p3d[c] = Xc + rng.normal(0, sigma_3d, (N, N_SLOTS, 3))

    if score_mode == "ones":
        s2d = np.ones((C, N, N_SLOTS))
    else:
        s2d = rng.uniform(0.2, 1.0, (C, N, N_SLOTS))
    s3d = np.zeros((C, N, N_SLOTS))
    s3d[:, :, :15] = 1.0          # what H36M-17 -> OpenPose-25 fills (reference inference.py:46-59)
    p3d[:, :, 15:, :] = 0.0

    if p_drop_2d > 0:
        drop = rng.random((C, N, N_SLOTS)) < p_drop_2d
        # keep >= 2 views per (frame, joint): un-drop the first two cameras where needed
        nvis = (~drop).sum(axis=0)
        need = nvis < 2
        drop[0][need] = False
        drop[1][need] = False
        s2d[drop] = 0.0
    if p_drop_3d > 0:
        drop3 = rng.random((C, N, N_SLOTS)) < p_drop_3d
        s3d[drop3] = 0.0

    R_init = np.zeros_like(R_gt)
    t_init = np.zeros_like(t_gt)
    for c in range(C):
        R_init[c] = rodrigues(rng.normal(0, init_rot_sigma, 3)) @ R_gt[c]
        t_init[c] = t_gt[c] + rng.normal(0, init_t_sigma, (3, 1))

    if camid is None:
        camid = np.array([21, 23, 25, 27]) if C == 4 else np.arange(C)
    return Scene(np.asarray(camid), K, R_gt, t_gt, R_init, t_init, X_w, p2d, s2d, p3d, s3d,
                 seed, width, height)


def ba_inputs(scene, available_joints=None, frame_skip=1):
    """Slice a scene the way ba.py's __main__ does (reference ba.py:317-322):
    2-D arrays indexed by available_joints, 3-D arrays keep all 25 slots."""
    if available_joints is None:
        available_joints = list(range(17))
    aj = np.asarray(available_joints, dtype=np.int64)
    sp2d = np.ascontiguousarray(scene.p2d[:, ::frame_skip][:, :, aj, :])
    ss2d = np.ascontiguousarray(scene.s2d[:, ::frame_skip][:, :, aj])
    sp3d = np.ascontiguousarray(scene.p3d[:, ::frame_skip])
    ss3d = np.ascontiguousarray(scene.s3d[:, ::frame_skip])
    return sp2d, ss2d, sp3d, ss3d


# BASELINE.json config shapes -------------------------------------------------
def cfg1(seed=0, N=1000):
    """4 cameras (C021/23/25/27), 17 joints, N frames, sigma = 3 px."""
    return make_scene(C=4, N=N, seed=seed, sigma_px=3.0)


def cfg3(seed=0, N=10000, C=31):
    """Panoptic-shaped: 31 HD cameras, partial visibility."""
    return make_scene(C=C, N=N, seed=seed, sigma_px=3.0, score_mode="uniform",
                      p_drop_2d=0.3, p_drop_3d=0.1, width=1920, height=1080, focal=1400.0)


def cfg4(seed=0, N=100000, C=64):
    return make_scene(C=C, N=N, seed=seed, sigma_px=3.0, width=1920, height=1080, focal=1400.0)
