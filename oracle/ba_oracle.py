"""CPU ORACLE (test infrastructure, not product code) for the bundle-adjustment path.

A NumPy restatement of the reference's BA objective and of the SciPy trust-region
driver it is minimised with.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product path
(the CUDA library behind include/bax.h) never does.

Parity status: PINNED by execution.  The reference holds no tests, fixtures or
golden vectors for this path; tests/golden/*.npz are outputs of the
reference's own functions run in the build container through oracle/ref_shim.py
(generator: oracle/gen_golden.py) and tests/test_oracle_golden.py checks every
function below against them.

Each function cites the reference lines it follows (paths relative to the
reference checkout).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp

SQRT2 = np.sqrt(2.0)

# util.py:129-145 (OP_BONE through OP_KEY, util.py:20-46): OpenPose-25 indices of the 12 bones
OP_BONE = np.array([[8, 1], [2, 5], [2, 3], [5, 6], [4, 3], [7, 6],
                    [9, 8], [12, 8], [9, 10], [10, 11], [12, 13], [13, 14]], dtype=np.int64)


# --------------------------------------------------------------------------- SO(3)
def hat(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0.0]])


def rodrigues(rvec):
    """rvec -> R.  cv2.Rodrigues forward map as used at ba.py:29 (textbook formula)."""
    r = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = np.sqrt(r @ r)
    if th < np.finfo(np.float64).eps:      # OpenCV: theta < DBL_EPSILON -> identity
        return np.eye(3)
    k = r / th
    c, s = np.cos(th), np.sin(th)
    return c * np.eye(3) + (1 - c) * np.outer(k, k) + s * hat(k)


def rodrigues_inv(R):
    """R -> rvec with OpenCV's semantics (ba.py:19): the input is first projected on
    SO(3) by SVD (R <- U Vt), then the log map is taken, including the theta~pi branch."""
    U, _, Vt = np.linalg.svd(np.asarray(R, dtype=np.float64))
    R = U @ Vt
    r = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    s = np.sqrt((r @ r) * 0.25)
    c = np.clip((np.trace(R) - 1) * 0.5, -1.0, 1.0)
    th = np.arccos(c)
    if s < 1e-5:
        if c > 0:
            return np.zeros(3)
        t = (R[0, 0] + 1) * 0.5
        rx = np.sqrt(max(t, 0.0))
        t = (R[1, 1] + 1) * 0.5
        ry = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 1] < 0 else 1.0)
        t = (R[2, 2] + 1) * 0.5
        rz = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 2] < 0 else 1.0)
        if abs(rx) < abs(ry) and abs(rx) < abs(rz) and (R[1, 2] > 0) != (ry * rz > 0):
            rz = -rz
        v = np.array([rx, ry, rz])
        return v * (th / np.linalg.norm(v))
    return r * (0.5 * th / s)


def to_theta(R, t, x):
    """ba.py:18-20: theta = [all rvecs | all tvecs | all points]."""
    rv = np.concatenate([rodrigues_inv(Rc) for Rc in R])
    return np.concatenate([rv, np.asarray(t).reshape(-1), np.asarray(x).reshape(-1)])


def from_theta(theta, C):
    """ba.py:23-31."""
    rv = theta[:3 * C].reshape(C, 3)
    t = theta[3 * C:6 * C].reshape(C, 3, 1)
    x = theta[6 * C:].reshape(-1, 3)
    return np.stack([rodrigues(r) for r in rv]), t, x


def d_rotated_d_rvec(rvec, x):
    """d(R(r) x)/dr, 3x3 (Gallego & Yezzi 2015).  SURVEY.md §3.3; limit -hat(x) at r=0
    (linear_*/ransac_*.json carry R_0 = I exactly, so rvec_0 = 0 does occur)."""
    r = np.asarray(rvec, dtype=np.float64)
    n2 = r @ r
    if n2 < np.finfo(np.float64).eps ** 2:
        return -hat(x)
    R = rodrigues(r)
    return -R @ hat(x) @ (np.outer(r, r) + (R.T - np.eye(3)) @ hat(r)) / n2


# --------------------------------------------------------------------------- residual families
def residual_reproj(K, R, t, x, y, s2d_flat):
    """ba.py:34-63 (objfun_nll): camera-major, visible points only, (ex, ey) interleaved,
    weight sqrt(2)*score.  Returns (r, list of visible point indices per camera)."""
    out, vis = [], []
    for c in range(len(K)):
        idx = np.nonzero(s2d_flat[c] > 0)[0]
        P = x[idx] @ R[c].T + t[c].reshape(1, 3)
        q = P @ K[c].T
        yh = q[:, :2] / q[:, 2:3]
        w = SQRT2 * s2d_flat[c, idx]
        out.append(((y[c, idx] - yh) * w[:, None]).reshape(-1))
        vis.append(idx)
    return np.concatenate(out), vis


def bone_directions(p3d, s3d, bones):
    """Constant data of the orientation term (ba.py:73-91): unit camera-frame bone
    directions for the (frame, bone) rows whose two endpoints have score > 0 in EVERY
    camera and whose direction is finite.  Returns (dhat (C, Nv, 3), rows (Nv,) = n*B+b)."""
    C, N = p3d.shape[:2]
    B = len(bones)
    a = p3d[:, :, bones[:, 0], :]
    b = p3d[:, :, bones[:, 1], :]
    ok = (s3d > 0)
    ok = ok[:, :, bones[:, 0]] & ok[:, :, bones[:, 1]]          # (C,N,B)
    with np.errstate(invalid="ignore", divide="ignore"):
        d = a - b
        d = d / np.linalg.norm(d, axis=-1, keepdims=True)
    ok &= ~np.isnan(d).any(axis=-1)
    valid = ok.all(axis=0).reshape(N * B)
    rows = np.nonzero(valid)[0]
    return d.reshape(C, N * B, 3)[:, rows, :], rows


def residual_var3d(R, dhat):
    """ba.py:67-96 (objfun_var3d): 1 - || mean_c R_c^T dhat_c || per valid (frame, bone)."""
    C = len(R)
    m = np.zeros(dhat.shape[1:])
    for c in range(C):
        m += dhat[c] @ R[c]          # row-vector form of R_c^T d
    m /= C
    return 1.0 - np.linalg.norm(m, axis=1)


def residual_varbone(x_NJ3, bones):
    """ba.py:99-110 (objfun_varbone): population variance over frames of each bone length."""
    dv = x_NJ3[:, bones[:, 0], :] - x_NJ3[:, bones[:, 1], :]
    L = np.linalg.norm(dv, axis=2)
    return np.var(L, axis=0)


def objfun(params, K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2):
    """ba.py:113-140, same signature: sp2d (C,N*J,2), ss2d (C,N,J), sp3d (C,N,25,3), ss3d (C,N,25)."""
    R, t, x = from_theta(np.asarray(params, dtype=np.float64), C)
    e1, _ = residual_reproj(K, R, t, x, sp2d.reshape(C, N * J, 2), ss2d.reshape(C, N * J))
    dhat, _ = bone_directions(sp3d, ss3d, bone_idx)
    e2 = residual_var3d(R, dhat) * lambda1
    e3 = residual_varbone(x.reshape(N, J, 3), bone_idx) * lambda2
    return np.concatenate([e1, e2, e3])


def jacobian(params, K, sp2d, ss2d, sp3d, ss3d, bone_idx, C, N, J, lambda1, lambda2):
    """Analytic Jacobian of `objfun` as scipy CSR (m x n).  The reference has none (it lets
    SciPy finite-difference ba.objfun, ba.py:187-206); formulas: SURVEY.md §3.3."""
    theta = np.asarray(params, dtype=np.float64)
    R, t, x = from_theta(theta, C)
    rv = theta[:3 * C].reshape(C, 3)
    y = sp2d.reshape(C, N * J, 2)
    s = ss2d.reshape(C, N * J)
    n_par = theta.size
    rows, cols, vals = [], [], []
    row0 = 0
    for c in range(C):
        idx = np.nonzero(s[c] > 0)[0]
        for k, p in enumerate(idx):
            P = R[c] @ x[p] + t[c, :, 0]
            q = K[c] @ P
            yh = q[:2] / q[2]
            w = SQRT2 * s[c, p]
            dedP = -(w / q[2]) * (np.array([[1, 0, -yh[0]], [0, 1, -yh[1]]]) @ K[c])
            Jr = dedP @ d_rotated_d_rvec(rv[c], x[p])
            Jx = dedP @ R[c]
            for a in range(2):
                r_ = row0 + 2 * k + a
                for b in range(3):
                    rows += [r_, r_, r_]
                    cols += [3 * c + b, 3 * C + 3 * c + b, 6 * C + 3 * p + b]
                    vals += [Jr[a, b], dedP[a, b], Jx[a, b]]
        row0 += 2 * len(idx)
    dhat, _ = bone_directions(sp3d, ss3d, bone_idx)
    Nv = dhat.shape[1]
    m_ = np.zeros((Nv, 3))
    for c in range(C):
        m_ += dhat[c] @ R[c]
    m_ /= C
    mh = m_ / np.linalg.norm(m_, axis=1, keepdims=True)
    for c in range(C):
        for i in range(Nv):
            # d(R(r)^T d)/dr = -dRx(-r, d)
            g = -(lambda1 / C) * (mh[i] @ (-d_rotated_d_rvec(-rv[c], dhat[c, i])))
            for b in range(3):
                rows.append(row0 + i); cols.append(3 * c + b); vals.append(g[b])
    row0 += Nv
    X = x.reshape(N, J, 3)
    dv = X[:, bone_idx[:, 0], :] - X[:, bone_idx[:, 1], :]
    L = np.linalg.norm(dv, axis=2)
    Lm = L.mean(axis=0)
    for b in range(len(bone_idx)):
        coef = lambda2 * (2.0 / N) * (L[:, b] - Lm[b]) / L[:, b]
        for n in range(N):
            gvec = coef[n] * dv[n, b]
            pa = n * J + bone_idx[b, 0]
            pb = n * J + bone_idx[b, 1]
            for k in range(3):
                rows += [row0 + b, row0 + b]
                cols += [6 * C + 3 * pa + k, 6 * C + 3 * pb + k]
                vals += [gvec[k], -gvec[k]]
    row0 += len(bone_idx)
    return sp.csr_matrix((vals, (rows, cols)), shape=(row0, n_par))


# --------------------------------------------------------------------------- initialisation
def triangulate_with_conf(p2d, s2d, K, R, t, mask):
    """util.py:325-354 (+ triangulate_point 308-322, constraint_mat 287-305): per (frame, joint)
    confidence-weighted linear triangulation; NaN unless sum(conf) > 0 and >= 2 masked views;
    NaN if <= 2 positive weighted rows; raises if rank < 3."""
    C, N, J, _ = p2d.shape
    P = np.stack([K[c] @ np.hstack([R[c], t[c].reshape(3, 1)]) for c in range(C)])
    X = np.full((N, J, 3), np.nan)
    for n in range(N):
        for j in range(J):
            m = mask[:, n, j]
            conf = s2d[:, n, j]
            if not (conf.sum() > 0 and m.sum() > 1):
                continue
            A, b, w = [], [], []
            for c in np.nonzero(m)[0]:
                u, v = p2d[c, n, j]
                A.append(u * P[c, 2, :3] - P[c, 0, :3]); b.append(-(u * P[c, 2, 3] - P[c, 0, 3]))
                A.append(v * P[c, 2, :3] - P[c, 1, :3]); b.append(-(v * P[c, 2, 3] - P[c, 1, 3]))
                w += [conf[c], conf[c]]
            A, b, w = np.array(A), np.array(b), np.array(w)
            sol, _, rank, _ = np.linalg.lstsq(A * w[:, None], b * w, rcond=None)
            if np.sum(w > 0) <= 2:
                continue
            if rank < 3:
                raise Exception("not enough constraint")
            X[n, j] = sol
    return X


# --------------------------------------------------------------------------- SciPy TRF (exact tr_solver) restated
EPS = np.finfo(np.float64).eps


def solve_lsq_trust_region(n, m, uf, s, V, Delta, initial_alpha=None, rtol=0.01, max_iter=10):
    """scipy/optimize/_lsq/common.py:57-168 (More' & Sorensen root-finding on the SVD of J)."""
    def phi_and_derivative(alpha):
        denom = s ** 2 + alpha
        p_norm = np.linalg.norm(suf / denom)
        return p_norm - Delta, -np.sum(suf ** 2 / denom ** 3) / p_norm

    suf = s * uf
    full_rank = (m >= n) and (s[-1] > EPS * m * s[0])
    if full_rank:
        p = -V.dot(uf / s)
        if np.linalg.norm(p) <= Delta:
            return p, 0.0, 0
    alpha_upper = np.linalg.norm(suf) / Delta
    if full_rank:
        phi, phi_prime = phi_and_derivative(0.0)
        alpha_lower = -phi / phi_prime
    else:
        alpha_lower = 0.0
    if initial_alpha is None or (not full_rank and initial_alpha == 0):
        alpha = max(0.001 * alpha_upper, (alpha_lower * alpha_upper) ** 0.5)
    else:
        alpha = initial_alpha
    for it in range(max_iter):
        if alpha < alpha_lower or alpha > alpha_upper:
            alpha = max(0.001 * alpha_upper, (alpha_lower * alpha_upper) ** 0.5)
        phi, phi_prime = phi_and_derivative(alpha)
        if phi < 0:
            alpha_upper = alpha
        ratio = phi / phi_prime
        alpha_lower = max(alpha_lower, alpha - ratio)
        alpha -= (phi + Delta) * ratio / Delta
        if np.abs(phi) < rtol * Delta:
            break
    p = -V.dot(suf / (s ** 2 + alpha))
    p *= Delta / np.linalg.norm(p)
    return p, alpha, it + 1


def solve_tr_floored(uf, s, V, Delta, initial_alpha, alpha_floor, rtol=0.01, max_iter=10):
    """The product's variant of the More' loop (package ba._solve_tr), restated on the SVD: rank-deficient
    branch of SciPy's solver with alpha floored at `alpha_floor`; if the step is still inside the region
    at the floor it is returned unscaled (interior step).  -> (p, alpha, n_iter, on_boundary)"""
    suf = s * uf

    def step(alpha):
        return -V.dot(suf / (s ** 2 + alpha))

    alpha_upper = np.linalg.norm(suf) / Delta
    alpha_lower = 0.0

    def reset():
        return max(0.001 * alpha_upper, (alpha_lower * alpha_upper) ** 0.5, alpha_floor)

    alpha = reset() if (initial_alpha is None or initial_alpha == 0) else max(initial_alpha, alpha_floor)
    for it in range(max_iter):
        if alpha < alpha_lower or alpha > alpha_upper:
            alpha = reset()
        alpha = max(alpha, alpha_floor)
        denom = s ** 2 + alpha
        p_norm = np.linalg.norm(suf / denom)
        phi = p_norm - Delta
        if phi < 0 and alpha <= alpha_floor:
            return step(alpha), alpha, it + 1, False
        phi_prime = -np.sum(suf ** 2 / denom ** 3) / p_norm
        if phi < 0:
            alpha_upper = alpha
        ratio = phi / phi_prime
        alpha_lower = max(alpha_lower, alpha - ratio)
        alpha -= (phi + Delta) * ratio / Delta
        if np.abs(phi) < rtol * Delta:
            break
    alpha = max(alpha, alpha_floor)
    p = step(alpha)
    return p * (Delta / np.linalg.norm(p)), alpha, it + 1, True


def update_tr_radius(Delta, actual_reduction, predicted_reduction, step_norm, bound_hit):
    """scipy/optimize/_lsq/common.py:222-248."""
    if predicted_reduction > 0:
        ratio = actual_reduction / predicted_reduction
    elif predicted_reduction == actual_reduction == 0:
        ratio = 1
    else:
        ratio = 0
    if ratio < 0.25:
        Delta = 0.25 * step_norm
    elif ratio > 0.75 and bound_hit:
        Delta *= 2.0
    return Delta, ratio


def check_termination(dF, F, dx_norm, x_norm, ratio, ftol, xtol):
    """scipy/optimize/_lsq/common.py:705-717."""
    ftol_satisfied = dF < ftol * F and ratio > 0.25
    xtol_satisfied = dx_norm < xtol * (xtol + x_norm)
    if ftol_satisfied and xtol_satisfied:
        return 4
    if ftol_satisfied:
        return 2
    if xtol_satisfied:
        return 3
    return None


def trf_exact(fun, jac, x0, ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None, floor_cols=None):
    """scipy/optimize/_lsq/trf.py:415-587 (trf_no_bounds, tr_solver='exact', x_scale=1) with a
    dense Jacobian callback.  Returns dict(x, cost, nfev, njev, status, history).
    floor_cols: None = SciPy's solver verbatim; an int k = the product's floored variant
    (solve_tr_floored with alpha_floor = 1e-13 * max_j>=k (J^T J)_jj, k = 6C: point columns)."""
    x = np.array(x0, dtype=np.float64)
    f = fun(x)
    nfev, njev = 1, 1
    J = jac(x)
    m, n = J.shape
    cost = 0.5 * f @ f
    g = J.T @ f
    Delta = np.linalg.norm(x)
    if Delta == 0:
        Delta = 1.0
    if max_nfev is None:
        max_nfev = 100 * n
    alpha = 0.0
    status = None
    hist = []
    while True:
        if np.linalg.norm(g, np.inf) < gtol:
            status = 1
        if status is not None or nfev == max_nfev:
            break
        U, s, Vt = np.linalg.svd(J, full_matrices=False)
        V = Vt.T
        uf = U.T @ f
        actual_reduction = -1.0
        while actual_reduction <= 0 and nfev < max_nfev:
            if floor_cols is None:
                step, alpha, n_it = solve_lsq_trust_region(n, m, uf, s, V, Delta, initial_alpha=alpha)
            else:
                floor = 1e-13 * np.max(np.sum(J[:, floor_cols:] ** 2, axis=0))
                step, alpha, n_it, on_b = solve_tr_floored(uf, s, V, Delta, alpha, floor)
            Js = J @ step
            predicted_reduction = -(0.5 * Js @ Js + g @ step)
            x_new = x + step
            f_new = fun(x_new)
            nfev += 1
            step_norm = np.linalg.norm(step)
            if not np.all(np.isfinite(f_new)):
                Delta = 0.25 * step_norm
                continue
            cost_new = 0.5 * f_new @ f_new
            actual_reduction = cost - cost_new
            Delta_new, ratio = update_tr_radius(Delta, actual_reduction, predicted_reduction,
                                                step_norm, step_norm > 0.95 * Delta)
            hist.append(dict(alpha=alpha, Delta=Delta, step_norm=step_norm, cost_new=cost_new,
                             pred=predicted_reduction, act=actual_reduction, n_it=n_it))
            status = check_termination(actual_reduction, cost, step_norm, np.linalg.norm(x),
                                       ratio, ftol, xtol)
            if status is not None:
                break
            if floor_cols is not None and not on_b:
                alpha = 0.0                      # what SciPy's solver returns for an interior step (common.py:113-116)
            alpha *= Delta / Delta_new
            Delta = Delta_new
        if actual_reduction > 0:
            x, f, cost = x_new, f_new, cost_new
            J = jac(x)
            njev += 1
            g = J.T @ f
    if status is None:
        status = 0
    return dict(x=x, cost=cost, nfev=nfev, njev=njev, status=status, history=hist,
                optimality=np.linalg.norm(g, np.inf))


def ba_main(camid, K, R_w2c, t_w2c, sp2d, ss2d, sp3d, ss3d, lambda1, lambda2,
            ftol=1e-4, xtol=1e-8, gtol=1e-8, max_nfev=None, return_info=False, floored=False):
    """ba.py:157-210 driven by the restated TRF with the analytic Jacobian (exact-J semantics:
    the 6 gauge singular values are exactly zero, so the More' loop always starts from
    alpha_lower = 0 — the only intended difference from the FD-noise-driven reference path,
    SURVEY.md §7.2-1)."""
    C = len(camid)
    N, J = sp2d.shape[1], sp2d.shape[2]
    x0 = triangulate_with_conf(sp2d, ss2d, K, R_w2c, t_w2c, ss2d > 0).reshape(N * J, 3)
    theta0 = to_theta(R_w2c, t_w2c, x0)
    if not np.all(np.isfinite(theta0)):
        raise ValueError("Initial guess is outside of provided bounds")
    args = (K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, OP_BONE, C, N, J, lambda1, lambda2)
    res = trf_exact(lambda th: objfun(th, *args), lambda th: jacobian(th, *args).toarray(),
                    theta0, ftol=ftol, xtol=xtol, gtol=gtol, max_nfev=max_nfev,
                    floor_cols=6 * C if floored else None)
    R, t, x = from_theta(res["x"], C)
    return (R, t, x, res) if return_info else (R, t, x)


# --------------------------------------------------------------------------- gauge-free comparison
def gauge_normalise(R, t):
    """Express cameras relative to camera 0 and scale so the 0-1 baseline is 1
    (SURVEY.md §7.2-1): the BA cost is invariant to a global similarity."""
    R = np.asarray(R); t = np.asarray(t).reshape(-1, 3)
    Rn = np.stack([Rc @ R[0].T for Rc in R])
    tn = np.stack([t[c] - Rn[c] @ t[0] for c in range(len(R))])
    sc = np.linalg.norm(tn[1])
    return Rn, tn / sc


def rot_angle(Ra, Rb):
    c = (np.trace(Ra.T @ Rb) - 1) / 2
    if c > 1 - 1e-8:                      # stable small-angle branch
        return np.linalg.norm(Ra - Rb) / np.sqrt(2)
    return np.arccos(np.clip(c, -1, 1))
