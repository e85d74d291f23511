"""The oracle (oracle/*.py) against golden vectors produced by the unmodified reference
(oracle/gen_golden.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import ba_oracle as bo
from oracle import ransac_oracle as ro

BA_CASES = ["small", "partial", "synadl26", "gafa_lambda"]


def _ba_args(g):
    C, N, J = g["ss2d"].shape
    return (g["K"], g["sp2d"].reshape(C, N * J, 2), g["ss2d"], g["sp3d"], g["ss3d"], bo.OP_BONE,
            C, N, J, float(g["lam1"]), float(g["lam2"]))


@pytest.mark.parametrize("case", BA_CASES)
def test_objfun_matches_reference(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ba_{case}.npz"))
    for th, f in ((g["theta0"], g["f0"]), (g["theta1"], g["f1"])):
        r = bo.objfun(th, *_ba_args(g))
        assert r.shape == f.shape
        # north_star: residual vectors within 1e-9 relative (FP64)
        assert np.max(np.abs(r - f)) <= 1e-9 * max(1.0, np.max(np.abs(f)))


@pytest.mark.parametrize("case", BA_CASES)
def test_analytic_jacobian_matches_reference_fd(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ba_{case}.npz"))
    Ja = bo.jacobian(g["theta1"], *_ba_args(g)).toarray()
    Jfd = g["J_fd"]                       # 3-point FD of the reference's ba.objfun
    assert Ja.shape == Jfd.shape
    assert np.max(np.abs(Ja - Jfd)) <= 2e-6 * np.max(np.abs(Jfd))


@pytest.mark.parametrize("case", BA_CASES)
def test_triangulation_and_theta(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ba_{case}.npz"))
    x0 = bo.triangulate_with_conf(g["sp2d"], g["ss2d"], g["K"], g["R_init"], g["t_init"], g["ss2d"] > 0)
    assert np.allclose(x0, g["x0"], rtol=0, atol=1e-10, equal_nan=True)
    th = bo.to_theta(g["R_init"], g["t_init"], x0.reshape(-1, 3))
    assert np.max(np.abs(th - g["theta0"])) < 1e-12


def test_trf_restatement_matches_scipy(golden_dir):
    """trf_exact == scipy least_squares(method='trf', tr_solver='exact') on the same callbacks."""
    from scipy.optimize import least_squares
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    args = _ba_args(g)
    fun = lambda th: bo.objfun(th, *args)
    jac = lambda th: bo.jacobian(th, *args).toarray()
    res = least_squares(fun, g["theta0"], jac=jac, method="trf", ftol=1e-4)
    mine = bo.trf_exact(fun, jac, g["theta0"], ftol=1e-4)
    assert mine["nfev"] == res.nfev and mine["status"] == res.status
    assert np.max(np.abs(mine["x"] - res.x)) < 1e-9


def test_converged_extrinsics_match_reference(golden_dir):
    """Oracle BA (analytic J, tight tolerances) vs the reference objective driven to convergence
    by SciPy (FD Jacobian).  Gate: 1e-6 rad / 1e-6 relative t after gauge normalisation."""
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    C = g["K"].shape[0]
    R, t, x, info = bo.ba_main(np.arange(C), g["K"], g["R_init"], g["t_init"], g["sp2d"], g["ss2d"],
                               g["sp3d"], g["ss3d"], float(g["lam1"]), float(g["lam2"]),
                               ftol=1e-14, xtol=1e-14, gtol=1e-14, max_nfev=300, return_info=True, floored=True)
    Rn, tn = bo.gauge_normalise(R, t)
    Rr, tr = bo.gauge_normalise(g["R_tight3"], g["t_tight3"])
    ang = max(bo.rot_angle(Rn[c], Rr[c]) for c in range(C))
    dt = np.max(np.linalg.norm(tn - tr, axis=1))
    assert ang < 1e-6 and dt < 1e-6, (ang, dt)


@pytest.mark.parametrize("case", ["c4", "c3"])
def test_ransac_oracle(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ransac_{case}.npz"))
    v, n = ro.ransac_inputs(g["p3d"], g["s3d"], g["p2d"], g["s2d"], g["K"], bo.OP_BONE)
    assert np.array_equal(v, g["vc"]) and np.array_equal(n, g["n"])
    (R, t, _), info = ro.calib_linear_ransac(v, n, g["K"], int(g["n_iter"]), float(g["th_2d"]),
                                             float(g["th_3d"]), int(g["seed"]), return_info=True)
    assert np.array_equal(info["sn"], g["sn"]) and np.array_equal(info["sm"], g["sm"])
    assert np.max(np.abs(info["scores"] - g["scores"]) / g["scores"]) < 1e-9
    assert info["best"] == int(g["best"])
    assert np.array_equal(info["mask_2d"], g["mask_2d"]) and np.array_equal(info["mask_3d"], g["mask_3d"])
    assert np.max(np.abs(R - g["R_fin"])) < 1e-9 and np.max(np.abs(t - g["t_fin"])) < 1e-9
    for h in range(int(g["n_iter"])):
        Rh, th_, _ = ro.calib_linear(v[:, g["sn"][h]], n[:, g["sm"][h]])
        assert np.max(np.abs(Rh - g["R_hyp"][h])) < 1e-9 and np.max(np.abs(th_ - g["t_hyp"][h])) < 1e-9
