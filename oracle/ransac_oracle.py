"""CPU ORACLE (test infrastructure, not product code) for the RANSAC linear-calibration path.

NumPy restatement of calib_ransac.py's model fit, per-hypothesis score, winner masks and
final inlier fit.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.

Parity status: PINNED by execution against the reference's own functions
(tests/golden/ransac_*.npz written by oracle/gen_golden.py; the reference has no tests or
fixtures of its own for this path).

Third-party arithmetic restated here because its source is not under the reference tree:
  * pycalib-simple 2022.9.2.2 `pycalib.calib.triangulate` (conda.yaml:159) — the reference
    vendors an identical copy at util.py:240-258, which `triangulate_dlt` follows;
  * OpenCV `cv2.triangulatePoints` (util.py:163-168): for two views, the right singular
    vector of the 4x4 DLT matrix belonging to the smallest singular value.
"""
from __future__ import annotations

import itertools

import numpy as np


def hat(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0.0]])


def rotations_from_directions(v_CxNx3):
    """calib_ransac.py:59-70: factorise the stacked bone directions; R_0 = I, the other
    R_c are NOT re-orthonormalised."""
    C = v_CxNx3.shape[0]
    V = np.hstack(list(v_CxNx3))                       # (Ns, 3C)
    _, _, Zt = np.linalg.svd(V)
    R_all = np.sqrt(C) * Zt[:3, :]
    R_all = np.linalg.inv(R_all[:3, :3]) @ R_all
    assert np.linalg.det(R_all[:3, :3]) > 0
    return R_all.T.reshape(-1, 3, 3)


def translation_system(R_list, n_CxMx3):
    """Dense form of the sparse system stacked at calib_ransac.py:72-91:
    collinearity rows (24-33) then coplanarity rows (36-47); unknowns [X_0..X_{M-1} | t_0..t_{C-1}]."""
    C, M = n_CxMx3.shape[:2]
    nu = 3 * (M + C)
    rows = []
    for c in range(C):
        for i in range(M):
            A = np.zeros((3, nu))
            nx = hat(n_CxMx3[c, i])
            A[:, 3 * i:3 * i + 3] = nx @ R_list[c]
            A[:, 3 * M + 3 * c:3 * M + 3 * c + 3] = nx
            rows.append(A)
    for a, b in itertools.combinations(range(C), 2):
        m = np.cross(n_CxMx3[a] @ R_list[a], n_CxMx3[b] @ R_list[b])      # (M,3)
        A = np.zeros((M, nu))
        A[:, 3 * M + 3 * a:3 * M + 3 * a + 3] = m @ R_list[a].T
        A[:, 3 * M + 3 * b:3 * M + 3 * b + 3] = -m @ R_list[b].T
        rows.append(A)
    return np.vstack(rows)


def triangulate_two_view(P1, P2, x1, x2):
    """cv2.triangulatePoints for one correspondence (see module docstring)."""
    A = np.stack([x1[0] * P1[2] - P1[0], x1[1] * P1[2] - P1[1],
                  x2[0] * P2[2] - P2[0], x2[1] * P2[2] - P2[1]])
    _, _, Vt = np.linalg.svd(A)
    X = Vt[3]
    return X[:3] / X[3]


def z_test_w2c(R1, t1, R2, t2, n1, n2):
    """util.py:161-180: triangulate with (+t) and (-t), count points in front of both cameras."""
    def tri(sgn):
        P1 = np.hstack([R1, sgn * t1[:, None]])
        P2 = np.hstack([R2, sgn * t2[:, None]])
        return np.array([triangulate_two_view(P1, P2, a[:2], b[:2]) for a, b in zip(n1, n2)])

    def zc(R, t, X):
        return int(np.sum((X @ R.T + t)[:, 2] > 0))

    Xp, Xn = tri(1.0), tri(-1.0)
    zp = zc(R1, t1, Xp) + zc(R2, t2, Xp)
    zn = zc(R1, t1, Xn) + zc(R2, t2, Xn)
    return (1 if zp > zn else -1), zp, zn


def calib_linear(v_CxNx3, n_CxMx3):
    """calib_ransac.py:51-131.  Returns (R (C,3,3), t (C,3,1), X (M,3))."""
    C, M = n_CxMx3.shape[:2]
    R_list = rotations_from_directions(v_CxNx3)
    A = translation_system(R_list, n_CxMx3)
    w, v = np.linalg.eigh(A.T @ A)
    k = v[:, :4]                                         # calib_ransac.py:96-107
    _, _, vt = np.linalg.svd(k[3 * M:3 * M + 3, :])      # rows of t_0
    u = k @ vt[3]
    X = u[:3 * M].reshape(-1, 3)
    t = u[3 * M:]
    s = np.linalg.norm(t[3:6])
    t = (t / s).reshape(-1, 3)
    X = X / s
    sign, _, _ = z_test_w2c(R_list[0], t[0], R_list[1], t[1], n_CxMx3[0], n_CxMx3[1])
    return R_list, (sign * t).reshape(-1, 3, 1), sign * X


def triangulate_dlt(x_Cx2, P_Cx3x4):
    """util.py:240-258: eigenvector of the smallest eigenvalue of sum_c a^T a."""
    AtA = np.zeros((4, 4))
    for c in range(len(P_Cx3x4)):
        a = np.stack([P_Cx3x4[c][0] - x_Cx2[c][0] * P_Cx3x4[c][2],
                      P_Cx3x4[c][1] - x_Cx2[c][1] * P_Cx3x4[c][2]])
        AtA += a.T @ a
    _, v = np.linalg.eigh(AtA)
    if np.isclose(v[3, 0], 0):
        return v[:, 0]
    return v[:, 0] / v[3, 0]


def score_hypothesis(R, t, v_CxNx3, n_CxMx3):
    """calib_ransac.py:154-180: returns (e (M,C,3) reprojection errors in normalised
    coordinates, e3d (N,) angular errors, e_sum, e3d_sum)."""
    C, M = n_CxMx3.shape[:2]
    N = v_CxNx3.shape[1]
    P = np.stack([np.hstack([R[c], t[c].reshape(3, 1)]) for c in range(C)])
    # all M DLT systems at once
    a0 = P[:, None, 0, :] - n_CxMx3[:, :, 0:1] * P[:, None, 2, :]      # (C,M,4)
    a1 = P[:, None, 1, :] - n_CxMx3[:, :, 1:2] * P[:, None, 2, :]
    AtA = np.einsum("cmi,cmj->mij", a0, a0) + np.einsum("cmi,cmj->mij", a1, a1)
    _, vecs = np.linalg.eigh(AtA)
    X = vecs[:, :, 0]                                    # scale / sign cancel in the projection below
    x2 = np.einsum("cij,mj->mci", P, X)
    x2 = x2 / x2[:, :, 2:3]
    e = x2 - np.transpose(n_CxMx3, (1, 0, 2))
    v0 = v_CxNx3[0] @ R[0]
    acc = np.zeros(N)
    for c in range(1, C):
        acc += np.einsum("ni,ni->n", v0, v_CxNx3[c] @ R[c])
    e3d = np.arccos(np.minimum(acc / (C - 1), 1.0))
    return e, e3d, np.linalg.norm(e.reshape(-1)) / M, np.linalg.norm(e3d) / N


def sample_indices(N, M, n_iter, seed):
    """The reference's sequential RNG stream (calib_ransac.py:148-151)."""
    rng = np.random.default_rng(seed)
    sn = np.zeros((n_iter, 10), dtype=np.int64)
    sm = np.zeros((n_iter, 10), dtype=np.int64)
    for it in range(n_iter):
        sn[it] = rng.choice(N, size=10, replace=False)
        sm[it] = rng.choice(M, size=10, replace=False)
    return sn, sm


def winner_masks(e, e3d, K, th_2d, th_3d):
    """calib_ransac.py:193-204."""
    d = np.linalg.norm(e[:, :, :2], axis=2).T            # (C,M)
    d = d * K[:, 0, 0][:, None]
    return np.min(d <= th_2d, axis=0), e3d <= th_3d


def calib_linear_ransac(v_CxNx3, n_CxMx3, K, n_iter, th_2d, th_3d, seed, *, final_fit=True,
                        return_info=False):
    """calib_ransac.py:134-209 (strict '<' keeps the first minimum; NaN scores never win)."""
    N, M = v_CxNx3.shape[1], n_CxMx3.shape[1]
    sn, sm = sample_indices(N, M, n_iter, seed)
    best, E_best, scores = -1, np.inf, np.zeros(n_iter)
    e_best = e3d_best = None
    for it in range(n_iter):
        R, t, _ = calib_linear(v_CxNx3[:, sn[it]], n_CxMx3[:, sm[it]])
        e, e3d, es, e3s = score_hypothesis(R, t, v_CxNx3, n_CxMx3)
        scores[it] = es + e3s
        if es + e3s < E_best:
            best, E_best, e_best, e3d_best = it, es + e3s, e, e3d
    mask_2d, mask_3d = winner_masks(e_best, e3d_best, K, th_2d, th_3d)
    out = calib_linear(v_CxNx3[:, mask_3d], n_CxMx3[:, mask_2d]) if final_fit else (None, None, None)
    if return_info:
        return out, dict(best=best, scores=scores, mask_2d=mask_2d, mask_3d=mask_3d, sn=sn, sm=sm)
    return out


# ---------------------------------------------------------------------------- pre-processing (r1)
def ransac_inputs(p3d, s3d, p2d, s2d, K, bones):
    """calib_ransac.py:236-270 + util.py:490-569: all-camera visibility mask, unit bone
    directions (e1 - e0), all-visible 2-D points (all 25 slots: joints2projections ignores
    its joint list, util.py:556-563) and K^-1 normalised rays."""
    C, N, Jn = s2d.shape
    vis = np.min((s2d > 0) * (s3d == 1), axis=0).astype(bool)          # (N,J)
    q = np.array(p3d, dtype=np.float64)
    q[:, ~vis, :] = np.nan
    d = q[:, :, bones[:, 1], :] - q[:, :, bones[:, 0], :]
    d = d.reshape(C, N * len(bones), 3)
    keep = ~np.isnan(d).any(axis=(0, 2))
    d = d[:, keep, :]
    d = d / np.linalg.norm(d, axis=2, keepdims=True)
    y = np.array(p2d, dtype=np.float64)
    y[:, ~vis, :] = np.nan
    y = y.reshape(C, -1, 2)
    y = y[:, ~np.isnan(y).any(axis=(0, 2)), :]
    n = np.ones((C, y.shape[1], 3))
    n[:, :, :2] = y
    n = np.stack([n[c] @ np.linalg.inv(K[c]).T for c in range(C)])
    return d, n
