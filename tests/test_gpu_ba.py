"""GPU parity tests of the BA path: CUDA kernels (through the C ABI, include/bax.h) vs the oracle and the
golden vectors produced by the unmodified reference.  Tolerances follow BASELINE.json's north_star:
residual vectors 1e-9 relative (FP64); converged extrinsics 1e-6 rad / 1e-6 relative translation."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import calib_b200  # noqa: E402,F401
from calib_b200 import ba, scene  # noqa: E402
from oracle import ba_oracle as bo  # noqa: E402

BA_CASES = ["small", "partial", "synadl26", "gafa_lambda"]


def _load(golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ba_{case}.npz"))
    C, N, J = g["ss2d"].shape
    args = (g["K"], g["sp2d"].reshape(C, N * J, 2), g["ss2d"], g["sp3d"], g["ss3d"], bo.OP_BONE,
            C, N, J, float(g["lam1"]), float(g["lam2"]))
    prob = ba.BAProblem(g["K"], g["sp2d"], g["ss2d"], g["sp3d"], g["ss3d"], ba.OP_BONE,
                        float(g["lam1"]), float(g["lam2"]))
    return g, args, prob


def _rel(a, b):
    return np.max(np.abs(a - b)) / max(1.0, np.max(np.abs(b)))


@pytest.mark.parametrize("case", BA_CASES)
def test_residual_vs_reference_golden(golden_dir, case):
    g, args, prob = _load(golden_dir, case)
    for th, f in ((g["theta0"], g["f0"]), (g["theta1"], g["f1"])):
        r = prob.residual(th).cpu().numpy()
        assert r.shape == f.shape
        assert _rel(r, f) <= 1e-9
        # drop-in function signature
        r2 = ba.objfun(th, *args)
        assert _rel(r2, f) <= 1e-9


@pytest.mark.parametrize("case", BA_CASES)
def test_jacobian_vs_oracle_and_reference_fd(golden_dir, case):
    g, args, prob = _load(golden_dir, case)
    Jg = prob.jacobian(g["theta1"]).toarray()
    Jo = bo.jacobian(g["theta1"], *args).toarray()
    assert Jg.shape == Jo.shape
    assert np.max(np.abs(Jg - Jo)) <= 1e-9 * np.max(np.abs(Jo))
    assert np.max(np.abs(Jg - g["J_fd"])) <= 2e-6 * np.max(np.abs(g["J_fd"]))


@pytest.mark.parametrize("case", BA_CASES)
def test_triangulation_vs_reference(golden_dir, case):
    g, args, prob = _load(golden_dir, case)
    x0 = prob.triangulate(g["R_init"], g["t_init"])
    assert np.allclose(x0, g["x0"], rtol=0, atol=1e-9, equal_nan=True)


@pytest.mark.parametrize("case", BA_CASES)
def test_cost_gradient_and_damped_step(golden_dir, case):
    g, args, prob = _load(golden_dir, case)
    th = g["theta1"]
    f = bo.objfun(th, *args)
    J = bo.jacobian(th, *args).toarray()
    grad = J.T @ f
    assert abs(prob.cost(th) - 0.5 * f @ f) <= 1e-11 * (0.5 * f @ f)
    cost, ginf, g2, xn = prob.linearize(th)
    assert abs(cost - 0.5 * f @ f) <= 1e-11 * cost
    assert abs(ginf - np.max(np.abs(grad))) <= 1e-9 * ginf
    assert abs(g2 - np.linalg.norm(grad)) <= 1e-9 * g2
    assert abs(xn - np.linalg.norm(th)) <= 1e-12 * xn
    H = J.T @ J
    for alpha in (1e-3 * g2 / xn, 1.0 * g2 / xn, 37.0):
        pn2, gp, ptq = prob.trial(alpha, True)
        A = H + alpha * np.eye(H.shape[0])
        p = np.linalg.solve(A, -grad)
        pg = prob.step.cpu().numpy()
        assert np.linalg.norm(pg - p) <= 1e-8 * np.linalg.norm(p), alpha
        assert abs(pn2 - p @ p) <= 1e-8 * (p @ p)
        assert abs(gp - grad @ p) <= 1e-8 * abs(grad @ p)
        q = np.linalg.solve(A, p)
        assert abs(ptq - p @ q) <= 1e-7 * abs(p @ q)


def test_tr_subproblem_vs_scipy_solver(golden_dir):
    """One trust-region sub-solve: the GPU More' loop (damped Schur solves) against SciPy's
    solve_lsq_trust_region restated on the SVD of the oracle Jacobian, same Delta and initial alpha."""
    g, args, prob = _load(golden_dir, "partial")
    th = g["theta1"]
    f = bo.objfun(th, *args)
    J = bo.jacobian(th, *args).toarray()
    U, s, Vt = np.linalg.svd(J, full_matrices=False)
    cost, ginf, g2, xn = prob.linearize(th)
    for Delta, a0 in ((0.05, 0.0), (0.3, 0.0), (0.3, 5.0)):
        p_ref, alpha_ref, nit_ref = bo.solve_lsq_trust_region(J.shape[1], J.shape[0], U.T @ f, s, Vt.T, Delta,
                                                              initial_alpha=a0)
        alpha, nit, p_norm, gdotp, on_b = ba._solve_tr(prob, Delta, g2, a0)
        assert on_b and nit == nit_ref
        assert abs(alpha - alpha_ref) <= 1e-7 * alpha_ref
        p = prob.step.cpu().numpy() * (Delta / p_norm)
        assert np.linalg.norm(p - p_ref) <= 1e-8 * Delta


def test_trf_follows_oracle_path(golden_dir):
    """Same TRF iteration sequence (alpha, Delta, cost per trial) as the oracle's restatement of SciPy's
    trf_no_bounds on the dense analytic Jacobian with the floored More' loop, literal ftol = 1e-4."""
    g, args, prob = _load(golden_dir, "small")
    C = g["K"].shape[0]
    ref = bo.trf_exact(lambda th: bo.objfun(th, *args), lambda th: bo.jacobian(th, *args).toarray(),
                       g["theta0"], ftol=1e-4, floor_cols=6 * C)
    res = ba.trf_schur(prob, g["theta0"], ftol=1e-4)
    assert res["nfev"] == ref["nfev"] and res["status"] == ref["status"]
    for a, b in zip(res["history"], ref["history"]):
        # interior steps at the alpha floor carry round-off-level gauge components (cost is invariant to them
        # only to first order), so trial costs agree to ~1e-6 and the sequences stay in lock-step
        assert abs(a["alpha"] - b["alpha"]) <= 1e-2 * b["alpha"]
        assert abs(a["Delta"] - b["Delta"]) <= 1e-3 * b["Delta"]
        assert abs(a["cost_new"] - b["cost_new"]) <= 1e-4 * b["cost_new"]
    Rg, tg, _ = ba.from_theta(res["x"].cpu().numpy(), C)
    Ro, to, _ = bo.from_theta(ref["x"], C)
    Rg, tg = bo.gauge_normalise(Rg, tg); Ro, to = bo.gauge_normalise(Ro, to)
    assert max(bo.rot_angle(Rg[c], Ro[c]) for c in range(C)) < 1e-4
    assert np.max(np.linalg.norm(tg - to, axis=1)) < 1e-4


def test_converged_extrinsics_vs_reference(golden_dir):
    """Final extrinsics within 1e-6 rad / 1e-6 relative translation of the reference's converged solution
    (reference objfun driven by SciPy with tight tolerances), after gauge normalisation."""
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    C = g["K"].shape[0]
    R, t, x, info = ba.ba_main(np.arange(C), g["K"], g["R_init"], g["t_init"], g["sp2d"], g["ss2d"],
                               g["sp3d"], g["ss3d"], float(g["lam1"]), float(g["lam2"]),
                               ftol=1e-14, xtol=1e-14, gtol=1e-14, max_nfev=300, verbose=False, return_info=True)
    Rn, tn = bo.gauge_normalise(R, t)
    Rr, tr = bo.gauge_normalise(g["R_tight3"], g["t_tight3"])
    ang = max(bo.rot_angle(Rn[c], Rr[c]) for c in range(C))
    dt = np.max(np.linalg.norm(tn - tr, axis=1))
    assert ang < 1e-6 and dt < 1e-6, (ang, dt)


TIGHT_CASES = ["partial_tight", "gafa_tight", "panoptic_tight", "n40_tight"]


@pytest.mark.parametrize("case", TIGHT_CASES)
def test_converged_extrinsics_vs_reference_more_scenes(golden_dir, case):
    """north_star gate 3 beyond the toy scene: partial visibility (lambda 3 / 7), GAFA weights
    (lambda2 = 1e5, run_gafa.sh:3), Panoptic weights (lambda1 = 100, run_panoptic.sh:2, HD intrinsics, dropped
    detections) and a 40-frame scene.  Reference = the reference's objfun under its own solver call (ba.py:187-206)
    with ftol = xtol = gtol = 1e-14 and 3-point differences (oracle/gen_golden.py gen_ba_tight); ours = ba_main with
    the same tolerances.  1e-6 rad / 1e-6 relative translation after gauge normalisation."""
    g = np.load(os.path.join(golden_dir, f"ba_{case}.npz"))
    C = g["K"].shape[0]
    R, t, x, info = ba.ba_main(np.arange(C), g["K"], g["R_init"], g["t_init"], g["sp2d"], g["ss2d"],
                               g["sp3d"], g["ss3d"], float(g["lam1"]), float(g["lam2"]),
                               ftol=1e-14, xtol=1e-14, gtol=1e-14, max_nfev=400, verbose=False, return_info=True)
    Rn, tn = bo.gauge_normalise(R, t)
    Rr, tr = bo.gauge_normalise(g["R_tight3"], g["t_tight3"])
    ang = max(bo.rot_angle(Rn[c], Rr[c]) for c in range(C))
    dt = np.max(np.linalg.norm(tn - tr, axis=1))
    assert ang < 1e-6 and dt < 1e-6, (ang, dt, info["cost"], float(g["cost_tight3"]))
    assert info["cost"] <= float(g["cost_tight3"]) * (1 + 1e-7)         # same minimum (a flat direction in the points remains)


def test_objfun_dropin_sees_in_place_mutation(golden_dir):
    """The drop-in objfun caches the device problem by CONTENT: mutating sp2d in place between calls (the reference's
    own joints2projections mutates its input, util.py:556-569) must change the answer exactly as for the reference."""
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    C, N, J = g["ss2d"].shape
    sp2d = np.array(g["sp2d"]).reshape(C, N * J, 2)
    args = [g["K"], sp2d, g["ss2d"], g["sp3d"], g["ss3d"], ba.OP_BONE, C, N, J, 1.0, 1.0]
    f_a = ba.objfun(g["theta0"], *args)
    assert _rel(f_a, g["f0"]) < 1e-9
    sp2d[1, 5, 0] += 7.0                                    # same array object, new content
    f_b = ba.objfun(g["theta0"], *args)
    f_o = bo.objfun(g["theta0"], *args)
    assert _rel(f_b, f_o) < 1e-9 and np.max(np.abs(f_b - f_a)) > 1.0


def test_ba_main_literal_vs_reference_literal(golden_dir, capsys):
    """Literal ftol=1e-4 run.  The reference's own literal end point moves by ~1e-4 rad when only its FD
    scheme changes, so this comparison is bounded by that path noise, not by 1e-6."""
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    C = g["K"].shape[0]
    R, t, x = ba.ba_main(np.arange(C), g["K"], g["R_init"], g["t_init"], g["sp2d"], g["ss2d"],
                         g["sp3d"], g["ss3d"], float(g["lam1"]), float(g["lam2"]))
    out = capsys.readouterr().out
    assert "Initial mean NLL:" in out and "termination condition is satisfied" in out
    Rn, tn = bo.gauge_normalise(R, t)
    Rr, tr = bo.gauge_normalise(g["R_lit"], g["t_lit"])
    ang = max(bo.rot_angle(Rn[c], Rr[c]) for c in range(C))
    dt = np.max(np.linalg.norm(tn - tr, axis=1))
    assert ang < 1e-3 and dt < 1e-3, (ang, dt)


def test_initial_prior_prints_are_unscaled_also_for_zero_weights(golden_dir, capsys):
    """ba_main prints the UNSCALED prior terms, as the reference does (ba.py:179-183: objfun_var3d / objfun_varbone
    evaluated directly) — also when a weight is 0 and the term cannot be recovered from the weighted residual
    (ADVICE r01)."""
    g = np.load(os.path.join(golden_dir, "ba_small.npz"))
    C = g["K"].shape[0]
    vals = {}
    for lam in ((1.0, 1.0), (0.0, 1.0), (1.0, 0.0)):
        ba.ba_main(np.arange(C), g["K"], g["R_init"], g["t_init"], g["sp2d"], g["ss2d"], g["sp3d"], g["ss3d"], *lam,
                   max_nfev=2)
        out = capsys.readouterr().out
        vals[lam] = [float(l.split(":")[1]) for l in out.splitlines() if l.startswith("Initial mean")]
        assert len(vals[lam]) == 3
    args = (g["K"], g["sp2d"].reshape(C, -1, 2), g["ss2d"], g["sp3d"], g["ss3d"], bo.OP_BONE, C, g["ss2d"].shape[1],
            g["ss2d"].shape[2], 1.0, 1.0)
    f = bo.objfun(g["theta0"], *args)
    n_vis = int((g["ss2d"] > 0).sum())
    B = len(bo.OP_BONE)
    ref = [np.mean(f[:2 * n_vis] ** 2), np.mean(f[2 * n_vis:-B] ** 2), np.mean(f[-B:] ** 2)]
    for lam, v in vals.items():
        assert np.allclose(v, ref, rtol=1e-9, atol=0), (lam, v, ref)


def test_nan_start_raises_like_reference():
    sc = scene.make_scene(C=4, N=4, seed=3)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    ss2d[:, 1, 2] = 0.0
    ss2d[0, 1, 2] = 1.0           # a single-view point -> NaN start (reference util.py:344-349)
    with pytest.raises(ValueError, match="Initial guess is outside of provided bounds"):
        ba.ba_main(sc.camid, sc.K, sc.R_init, sc.t_init, sp2d, ss2d, sp3d, ss3d, 1.0, 1.0, verbose=False)


@pytest.mark.parametrize("C,N,drop", [(4, 300, 0.0), (7, 120, 0.3), (31, 40, 0.3)])
def test_medium_scene_vs_oracle(C, N, drop):
    sc = scene.make_scene(C=C, N=N, seed=100 + C, score_mode="uniform", p_drop_2d=drop, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    J = ss2d.shape[2]
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 2.0, 50.0)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 2.0, 50.0)
    f = bo.objfun(th, *args)
    assert _rel(prob.residual(th).cpu().numpy(), f) <= 1e-9
    # fused residual+Jacobian kernel: same residuals, and J acts like the directional derivative
    r2, Jc, Jp, oc, op = prob.residual_jacobian(th)
    nv = prob.n_vis
    assert _rel(r2.cpu().numpy().reshape(-1), f[:2 * nv]) <= 1e-9
    rng = np.random.default_rng(0)
    d = rng.normal(size=th.shape)
    eps = 1e-6
    fd = (bo.objfun(th + eps * d, *args) - bo.objfun(th - eps * d, *args)) / (2 * eps)
    Jd = prob.jacobian(th) @ d
    assert np.max(np.abs(Jd - fd)) <= 1e-5 * np.max(np.abs(fd))


def test_large_scene_properties():
    """BASELINE cfg1 size (C=4, N=1000, J=17): size-independent properties instead of an oracle run —
    cost == 0.5 |r|^2, the gradient equals J^T r, and a damped step solves its normal equations."""
    sc = scene.cfg1(seed=0)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 1.0)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    r = prob.residual(th).cpu().numpy()
    assert r.shape == (2 * 4 * 17000 + 12000 + 12,)
    J = prob.jacobian(th)
    cost, ginf, g2, xn = prob.linearize(th)
    grad = J.T @ r
    assert abs(cost - 0.5 * r @ r) <= 1e-11 * cost
    assert abs(g2 - np.linalg.norm(grad)) <= 1e-9 * g2
    alpha = 0.01 * g2 / xn
    pn2, gp, ptq = prob.trial(alpha, True)
    p = prob.step.cpu().numpy()
    resid = J.T @ (J @ p) + alpha * p + grad
    assert np.linalg.norm(resid) <= 1e-8 * np.linalg.norm(grad)
    assert abs(gp - grad @ p) <= 1e-9 * abs(gp)


# ----------------------------------------------------------------------------- paths added after the first GPU pass
@pytest.mark.parametrize("case", ["small", "partial", "gafa_lambda"])
def test_speculative_linearisation_same_path(golden_dir, case):
    """Linearising the trial point straight away (push / pop of the linearisation state) must walk exactly the
    iteration sequence of the cost-then-relinearise flow, rejected steps included."""
    g, args, prob = _load(golden_dir, case)
    hist = []
    for spec in (False, True):
        sol = ba.TRFSolver(prob, g["theta0"], ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=25, speculative=spec)
        while not sol.done():
            sol.iterate()
        hist.append((sol.history, sol.x.cpu().numpy().copy(), sol.nfev))
    (h0, x0, n0), (h1, x1, n1) = hist
    assert n0 == n1 and len(h0) == len(h1)
    for a, b in zip(h0, h1):
        assert abs(a["alpha"] - b["alpha"]) <= 1e-9 * abs(a["alpha"])
        assert abs(a["cost_new"] - b["cost_new"]) <= 1e-12 * abs(a["cost_new"])
    assert np.max(np.abs(x0 - x1)) <= 1e-9 * np.max(np.abs(x0))


def test_rejected_step_restores_linearisation(golden_dir):
    """A deliberately rejected trial (pop) leaves the damped solve of the old linearisation unchanged."""
    g, args, prob = _load(golden_dir, "partial")
    th = prob._dev_theta(g["theta0"])
    cost, ginf, g2, xn = prob.linearize(th)
    alpha = 0.01 * g2 / xn
    prob.trial(alpha, True)
    p0 = prob.step.clone()
    bad = th + 50.0 * torch.randn_like(th)             # a point far away: certainly a worse cost
    prob.linearize_push()
    cost_bad = prob.linearize(bad)[0]
    assert cost_bad > cost
    prob.linearize_pop()
    prob.trial(alpha, True)
    assert torch.equal(prob.step, p0)


@pytest.mark.parametrize("C,N", [(9, 60), (11, 40), (16, 30), (33, 20), (70, 6)])
def test_damped_step_many_cameras(C, N):
    """Reduced systems of 54..420 unknowns: DMMA SYRK tiles with the trailing rows fused into diagonal tiles (one to
    five tile rows, i.e. up to ten off-diagonal tiles and the per-kind K cut) and the cooperative multi-CTA reduced
    solve, against a dense solve of the oracle's normal equations."""
    sc = scene.make_scene(C=C, N=N, seed=300 + C, score_mode="uniform", p_drop_2d=0.2, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    J = ss2d.shape[2]
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 2.0, 30.0)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 2.0, 30.0)
    f = bo.objfun(th, *args)
    Jd = bo.jacobian(th, *args).toarray()
    cost, ginf, g2, xn = prob.linearize(th)
    grad = Jd.T @ f
    assert abs(g2 - np.linalg.norm(grad)) <= 1e-9 * g2
    for alpha in (1e-3 * g2 / xn, 1e-7 * g2 / xn):
        pn2, gp, ptq = prob.trial(alpha, True)
        A = Jd.T @ Jd + alpha * np.eye(Jd.shape[1])
        p = np.linalg.solve(A, -grad)
        pg = prob.step.cpu().numpy()
        assert np.linalg.norm(pg - p) <= 1e-8 * np.linalg.norm(p)
        assert abs(ptq - p @ np.linalg.solve(A, p)) <= 1e-7 * abs(ptq)


def test_single_cta_and_cooperative_reduced_solve_agree(monkeypatch):
    """BAX_DENSE_COOP=0 selects the single-CTA Cholesky; both must give the same step."""
    sc = scene.make_scene(C=12, N=40, seed=77, score_mode="uniform", p_drop_2d=0.2, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    steps = []
    for flag in ("0", "1"):
        monkeypatch.setenv("BAX_DENSE_COOP", flag)
        prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 10.0)
        x0 = prob.triangulate(sc.R_init, sc.t_init)
        th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        cost, ginf, g2, xn = prob.linearize(th)
        prob.trial(1e-4 * g2 / xn, True)
        steps.append(prob.step.cpu().numpy().copy())
        prob.close()
    assert np.linalg.norm(steps[0] - steps[1]) <= 1e-9 * np.linalg.norm(steps[0])


def test_observation_rebind(golden_dir):
    """bax_observations_rebind: a second device copy of the observations gives bit-identical residuals."""
    g, args, prob = _load(golden_dir, "partial")
    r0 = prob.residual(g["theta0"]).clone()
    y2, s2 = prob.y.clone(), prob.s.clone()
    prob.rebind_observations(y2, s2)
    assert torch.equal(prob.residual(g["theta0"]), r0)


@pytest.mark.parametrize("N", [37, 38, 15, 256])
@pytest.mark.parametrize("variant", ["1", "2", "3", "4", "5", "6", "7", "8", None])
def test_rj_kernel_variants_agree(monkeypatch, variant, N):
    """Every strategy of the fused residual + Jacobian kernel (LSU-staged, the TMA bulk-store variants, the persistent
    TMA-pipelined variants and the default choice) returns bit for bit the blocks of the plain LSU kernel, ragged
    visibility included.  N = 37: P = 629 is odd, the pipeline's 16-byte alignment rule sends it to the bulk-store
    kernel; N = 38: P = 646 = two full 256-tiles + a 134-observation tail; N = 15: a single partial tile; N = 256:
    full tiles only (P = 4352 = 17 x 256)."""
    sc = scene.make_scene(C=5, N=N, seed=41, score_mode="uniform", p_drop_2d=0.35, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    outs = []
    for v in ("0", variant):
        if v is None:
            monkeypatch.delenv("BAX_RJ_VARIANT", raising=False)
        else:
            monkeypatch.setenv("BAX_RJ_VARIANT", v)
        prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 1.0)
        x0 = prob.triangulate(sc.R_init, sc.t_init)
        th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        outs.append([t.cpu().numpy().copy() for t in prob.residual_jacobian(th)])
        prob.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


def test_rj_pipeline_all_invisible_tiles_and_nan_scores():
    """Tiles without a single visible observation (a camera that sees nothing for 300 frames), NaN scores and a
    camera with no observation at all go through the pipelined kernel and match the oracle's Jacobian blocks."""
    sc = scene.make_scene(C=4, N=400, seed=43, score_mode="uniform", p_drop_2d=0.1, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    ss2d = ss2d.copy()
    ss2d[1, 50:350] = 0.0                     # > 19 consecutive all-invisible 256-tiles in camera 1
    ss2d[2, ::7, 3] = np.nan                  # NaN score == invisible (ss2d > 0 is False)
    ss2d[3] = 0.0                             # camera 3 sees nothing
    C, N, J = ss2d.shape
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 1.0)
    x0 = np.nan_to_num(sc.X_w[:, :J].reshape(-1, 3))
    th = ba.to_theta(sc.R_init, sc.t_init, x0)
    r2, Jc, Jp, oc, op = [t.cpu().numpy() for t in prob.residual_jacobian(th)]
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 1.0, 1.0)
    f = bo.objfun(th, *args)
    n_vis = int((ss2d > 0).sum())
    assert r2.shape[0] == n_vis and not np.any(oc == 3)
    assert np.max(np.abs(r2.reshape(-1) - f[:2 * n_vis])) < 1e-9 * np.max(np.abs(f))
    Jd = bo.jacobian(th, *args).tocsr()[:2 * n_vis]
    rows = np.arange(n_vis)
    for a in range(2):
        for k in range(6):
            col = (3 * oc + k) if k < 3 else (3 * C + 3 * oc + (k - 3))
            ref = np.asarray(Jd[2 * rows + a, col]).reshape(-1)
            assert np.max(np.abs(Jc[:, a, k] - ref)) < 1e-9 * max(1.0, np.max(np.abs(ref)))
        for k in range(3):
            ref = np.asarray(Jd[2 * rows + a, 6 * C + 3 * op + k]).reshape(-1)
            assert np.max(np.abs(Jp[:, a, k] - ref)) < 1e-9 * max(1.0, np.max(np.abs(ref)))
    prob.close()


def test_ba_batch_equals_single_problem_runs():
    """BASELINE configs[4] (batch of independent 4-camera problems): the multi-stream sweep returns, problem by
    problem, bit for bit what a single ba_main call returns (same kernels, same order of operations), whatever
    the number of worker streams; a failing problem is reported with its index."""
    probs = []
    for k in range(7):
        sc = scene.make_scene(C=4, N=20 + 3 * k, seed=100 + k, score_mode="uniform", p_drop_2d=0.15, p_drop_3d=0.1)
        sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
        probs.append((sc.camid, sc.K, sc.R_init, sc.t_init, sp2d, ss2d, sp3d, ss3d))
    single = [ba.ba_main(*p, 1.0, 10.0, verbose=False) for p in probs]
    for n_streams in (1, 4):
        batch = ba.ba_batch(probs, 1.0, 10.0, n_streams=n_streams)
        assert len(batch) == len(single)
        for (R0, t0, x0), (R1, t1, x1) in zip(single, batch):
            assert np.array_equal(R0, R1) and np.array_equal(t0, t1) and np.array_equal(x0, x1)
    bad = list(probs[2])
    bad[5] = np.zeros_like(bad[5])                  # nothing visible: NaN start, as in the reference
    with pytest.raises(ValueError, match="problem 1"):
        ba.ba_batch([probs[0], tuple(bad), probs[1]], 1.0, 10.0, n_streams=2)


@pytest.mark.parametrize("which", ["cfg3", "cfg4shard"])
def test_full_size_configs_vs_oracle_and_step_identity(which):
    """BASELINE configs[2] at full size (31 cameras x 10 000 frames x 17 joints, 30 % dropped detections: 3.7e6
    observations, 510 186 parameters) and one GPU's shard of configs[3] (64 cameras x 12 500 frames: 1.36e7
    observations).  One oracle evaluation of ba.objfun is affordable at these sizes (seconds): residual vector
    1e-9.  The damped Schur step is checked through an identity that needs no assembled Jacobian: with
    (J^T J + alpha I) p = -g,  |J p|^2 = -g.p - alpha |p|^2,  J p taken as a central difference of the GPU
    residual along p."""
    sc = scene.cfg3(seed=5) if which == "cfg3" else scene.cfg4(seed=5, N=12500, C=64)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    C, N, J = ss2d.shape
    lam1, lam2 = (100.0, 1.0) if which == "cfg3" else (1.0, 1.0)
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, lam1, lam2)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    assert np.all(np.isfinite(th))
    r = prob.residual(th)
    f = bo.objfun(th, sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, lam1, lam2)
    assert r.shape[0] == f.shape[0] == 2 * prob.n_vis + prob.n_v3d + 12
    assert _rel(r.cpu().numpy(), f) <= 1e-9
    cost, ginf, g2, xn = prob.linearize(th)
    assert abs(cost - 0.5 * float(f @ f)) <= 1e-11 * cost
    alpha = 1e-3 * g2 / xn
    pn2, gp, ptq = prob.trial(alpha, True)
    assert gp < 0 and pn2 > 0 and ptq > 0
    p = prob.step.clone()
    thd = prob._dev_theta(th)
    eps = 1e-4 / float(p.abs().max())
    Jp = (prob.residual(thd + eps * p).clone() - prob.residual(thd - eps * p)) / (2 * eps)
    lhs = float(Jp @ Jp)
    rhs = -gp - alpha * pn2
    assert abs(lhs - rhs) <= 1e-6 * rhs
    # first-order decrease: g.p is the directional derivative of the cost along p
    dcost = (prob.cost(thd + eps * p) - prob.cost(thd - eps * p)) / (2 * eps)
    assert abs(dcost - gp) <= 1e-6 * abs(gp)
    prob.close()


@pytest.mark.parametrize("case", ["small", "partial", "gafa_lambda"])
def test_native_trust_region_driver_same_path(golden_dir, case):
    """The C-side driver (bax_trf_begin / _iterate / _solve, csrc/bax_trf.cu) and the Python statements of TRFSolver
    are the same algorithm: identical (alpha, Delta, step norm, cost, predicted, actual, More' iterations) sequence,
    rejected steps included, identical end point; bax_trf_solve in one call lands on the same point."""
    g, args, prob = _load(golden_dir, case)
    runs = []
    for native in (False, True):
        sol = ba.TRFSolver(prob, g["theta0"], ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=25, native=native)
        assert sol.native == native
        while not sol.done():
            sol.iterate()
        runs.append((sol.history, sol.x.cpu().numpy().copy(), sol.nfev, sol.njev, sol.status, sol.cost))
    (h0, x0, nf0, nj0, s0, c0), (h1, x1, nf1, nj1, s1, c1) = runs
    assert (nf0, nj0, s0) == (nf1, nj1, s1) and len(h0) == len(h1) and len(h0) >= 3
    for a, b in zip(h0, h1):
        for k in ("alpha", "Delta", "step_norm", "cost_new", "pred", "act"):
            assert abs(a[k] - b[k]) <= 1e-12 * max(abs(a[k]), 1e-300), (k, a[k], b[k])
        assert a["n_it"] == b["n_it"]
    assert np.max(np.abs(x0 - x1)) <= 1e-13 * np.max(np.abs(x0)) and abs(c0 - c1) <= 1e-13 * c0
    sol = ba.TRFSolver(prob, g["theta0"], ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=25, native=True)
    sol.solve_native()
    assert sol.done() and sol.nfev == nf1 and np.array_equal(sol.x.cpu().numpy(), x1)
    res = sol.result()
    assert res["status"] == (0 if s1 is None else s1) and res["njev"] == nj1
    prob.close()


@pytest.mark.parametrize("C,N,lam1,lam2,drop3d", [(2, 6, 1.0, 10.0, 0.0),      # the smallest rig the objective accepts
                                                  (4, 1, 1.0, 10.0, 0.0),      # one frame: bone-length variance is 0
                                                  (3, 9, 0.0, 0.0, 0.1),       # both priors switched off
                                                  (5, 7, 3.0, 0.0, 1.0),       # no 3-D pose valid anywhere: no orientation rows
                                                  (3, 40, 1.0, 1e5, 0.05)])    # GAFA-like lambda2 (run_gafa.sh:3)
def test_edge_shapes_vs_oracle(C, N, lam1, lam2, drop3d):
    """Ragged / degenerate shapes: residual vector, cost, gradient and one damped step against the oracle."""
    sc = scene.make_scene(C=C, N=N, seed=900 + C + N, score_mode="uniform", p_drop_2d=0.2 if C > 2 else 0.0,
                          p_drop_3d=min(drop3d, 0.99))
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    if drop3d >= 1.0:
        ss3d = np.zeros_like(ss3d)
    J = ss2d.shape[2]
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, lam1, lam2)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    assert np.all(np.isfinite(th))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, lam1, lam2)
    f = bo.objfun(th, *args)
    r = prob.residual(th).cpu().numpy()
    assert r.shape == f.shape and _rel(r, f) <= 1e-9
    if drop3d >= 1.0:
        assert prob.n_v3d == 0
    Jd = bo.jacobian(th, *args).toarray()
    cost, ginf, g2, xn = prob.linearize(th)
    grad = Jd.T @ f
    assert abs(cost - 0.5 * f @ f) <= 1e-10 * max(cost, 1e-300)
    assert abs(g2 - np.linalg.norm(grad)) <= 1e-8 * g2 and abs(ginf - np.max(np.abs(grad))) <= 1e-8 * ginf
    alpha = 0.01 * g2 / xn
    prob.trial(alpha, True)
    p_ref = np.linalg.solve(Jd.T @ Jd + alpha * np.eye(Jd.shape[1]), -grad)
    assert np.linalg.norm(prob.step.cpu().numpy() - p_ref) <= 1e-7 * np.linalg.norm(p_ref)
    R, t, x = ba.ba_main(sc.camid, sc.K, sc.R_init, sc.t_init, sp2d, ss2d, sp3d, ss3d, lam1, lam2, verbose=False)
    assert np.all(np.isfinite(R)) and np.all(np.isfinite(t)) and x.shape == (N * J, 3)
    prob.close()
