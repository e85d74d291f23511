"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

    python oracle/gen_golden.py            # needs /root/reference (or $CALIB_REFERENCE_DIR)

Every file stores its inputs next to the reference's outputs so the GPU box (which has no
reference checkout) can replay them.  Library versions are recorded in each file.
"""
import contextlib
import importlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

scene_mod = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200.scene")
OUT = os.path.join(ROOT, "tests", "golden")


def versions():
    import cv2
    import scipy
    return dict(numpy=np.__version__, scipy=scipy.__version__, cv2=cv2.__version__)


def quiet():
    return contextlib.redirect_stdout(io.StringIO())


def gen_ba(name, scene_kw, available_joints, lam1, lam2, run_main=False):
    rb, _, ru = ref_shim.load()
    from scipy.optimize import least_squares
    from scipy.optimize._numdiff import approx_derivative
    sc = scene_mod.make_scene(**scene_kw)
    sp2d, ss2d, sp3d, ss3d = scene_mod.ba_inputs(sc, available_joints)
    C, N, J = ss2d.shape
    x0 = ru.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init, ss2d > 0)
    theta0 = rb.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, ru.OP_BONE, C, N, J, lam1, lam2)
    f0 = rb.objfun(theta0, *args)
    # a second evaluation point away from the triangulated start (rvec_0 forced to exactly 0:
    # linear_*/ransac_*.json carry R_0 = I)
    rng = np.random.default_rng(123)
    theta1 = theta0 + rng.normal(0, 0.01, theta0.shape)
    theta1[0:3] = 0.0
    f1 = rb.objfun(theta1, *args)
    J_fd = approx_derivative(lambda th: rb.objfun(th, *args), theta1, method="3-point", abs_step=1e-6)
    out = dict(K=sc.K, R_init=sc.R_init, t_init=sc.t_init, sp2d=sp2d, ss2d=ss2d, sp3d=sp3d, ss3d=ss3d,
               lam1=lam1, lam2=lam2, x0=x0, theta0=theta0, f0=f0, theta1=theta1, f1=f1, J_fd=J_fd,
               versions=str(versions()))
    if run_main:
        with quiet():
            R, t, x = rb.ba_main(sc.camid, sc.K, sc.R_init, sc.t_init, sp2d, ss2d, sp3d, ss3d, lam1, lam2)
        out.update(R_lit=R, t_lit=t, x_lit=x)
        # the reference objective driven to convergence (same objfun, same solver, tight tolerances)
        for tag, jac in (("tight2", "2-point"), ("tight3", "3-point")):
            res = least_squares(rb.objfun, theta0, method="trf", jac=jac, ftol=1e-14, xtol=1e-14,
                                gtol=1e-14, max_nfev=300, args=args)
            R, t, x = rb.from_theta(res.x, C)
            out.update({f"R_{tag}": R, f"t_{tag}": t, f"cost_{tag}": res.cost, f"nfev_{tag}": res.nfev})
    np.savez_compressed(os.path.join(OUT, f"ba_{name}.npz"), **out)
    print("wrote", name, "m =", f0.size, "n =", theta0.size)


def gen_ba_tight(name, scene_kw, available_joints, lam1, lam2, max_nfev=300, jacs=("2-point", "3-point")):
    """Converged-extrinsics golden: the reference's objfun under the reference's solver call
    (ba.py:187-206) with ftol = xtol = gtol = 1e-14, plus the literal ba_main end point.  Stores inputs and end points
    only (no dense FD Jacobian), so a 40-frame scene stays a small file."""
    rb, _, ru = ref_shim.load()
    from scipy.optimize import least_squares
    sc = scene_mod.make_scene(**scene_kw)
    sp2d, ss2d, sp3d, ss3d = scene_mod.ba_inputs(sc, available_joints)
    C, N, J = ss2d.shape
    x0 = ru.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init, ss2d > 0)
    theta0 = rb.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, ru.OP_BONE, C, N, J, lam1, lam2)
    f0 = rb.objfun(theta0, *args)
    out = dict(K=sc.K, R_init=sc.R_init, t_init=sc.t_init, sp2d=sp2d, ss2d=ss2d, sp3d=sp3d, ss3d=ss3d,
               lam1=lam1, lam2=lam2, x0=x0, theta0=theta0, f0=f0, scene_kw=str(scene_kw), versions=str(versions()))
    with quiet():
        R, t, x = rb.ba_main(sc.camid, sc.K, sc.R_init, sc.t_init, sp2d, ss2d, sp3d, ss3d, lam1, lam2)
    out.update(R_lit=R, t_lit=t)
    for jac in jacs:
        tag = "tight2" if jac == "2-point" else "tight3"
        res = least_squares(rb.objfun, theta0, method="trf", jac=jac, ftol=1e-14, xtol=1e-14, gtol=1e-14,
                            max_nfev=max_nfev, args=args)
        R, t, x = rb.from_theta(res.x, C)
        out.update({f"R_{tag}": R, f"t_{tag}": t, f"cost_{tag}": res.cost, f"nfev_{tag}": res.nfev,
                    f"optimality_{tag}": res.optimality, f"theta_{tag}": res.x})
        print(name, tag, "nfev", res.nfev, "cost", res.cost, "optimality", res.optimality, flush=True)
    np.savez_compressed(os.path.join(OUT, f"ba_{name}.npz"), **out)
    print("wrote", name, "m =", f0.size, "n =", theta0.size, flush=True)


TIGHT_CASES = {
    # name: (scene_kw, joints, lambda1, lambda2)   lambdas: run_synadl.sh / run_gafa.sh:3 / run_panoptic.sh:2
    "partial_tight": (dict(C=5, N=8, seed=12, score_mode="uniform", p_drop_2d=0.3, p_drop_3d=0.1), 17, 3.0, 7.0),
    "gafa_tight": (dict(C=3, N=7, seed=14, score_mode="uniform"), 17, 1.0, 1e5),
    "panoptic_tight": (dict(C=6, N=8, seed=15, score_mode="uniform", p_drop_2d=0.3, p_drop_3d=0.1, width=1920,
                            height=1080, focal=1400.0), 17, 100.0, 1.0),
    "n40_tight": (dict(C=4, N=40, seed=16), 17, 1.0, 1.0),
}


def gen_ransac(name, scene_kw, n_iter, seed, th_2d, th_3d_deg):
    _, rc, ru = ref_shim.load()
    sc = scene_mod.make_scene(**scene_kw)
    C = sc.C
    mask = (sc.s2d > 0) * (sc.s3d == 1)
    vis = ru.visible_from_all_cam(mask)
    vc = ru.joints2orientations(sc.p3d, vis, ru.OP_BONE)
    y = ru.joints2projections(np.copy(sc.p2d), vis, ru.OP_KEY_SUB)
    n = np.ones((C, y.shape[1], 3))
    n[:, :, :2] = y
    n = np.array([n[i] @ np.linalg.inv(sc.K[i]).T for i in range(C)])
    th_3d = th_3d_deg / 180 * np.pi
    N, M = vc.shape[1], n.shape[1]

    # literal per-hypothesis pass: the reference's own calib_linear + triangulate, scored as at
    # calib_ransac.py:154-180
    rng = np.random.default_rng(seed)
    scores, Rs, ts, sn, sm = [], [], [], [], []
    for it in range(n_iter):
        a = rng.choice(N, size=10, replace=False)
        b = rng.choice(M, size=10, replace=False)
        R, t, _ = rc.calib_linear(vc[:, a, :], n[:, b, :])
        P = np.array([np.hstack((R[i], t[i].reshape((3, 1)))) for i in range(C)])
        e = []
        for i in range(M):
            X = rc.triangulate(n[:, i, :2], P)
            for c in range(C):
                x2 = P[c] @ X.reshape((4, 1))
                x2 = x2 / x2[2]
                e.append(x2.flatten() - n[c, i, :].flatten())
        e3d = []
        for i in range(N):
            e3 = 0
            for c in range(1, C):
                e3 += np.dot(vc[0, i] @ R[0], vc[c, i] @ R[c])
            e3d.append(np.min([e3 / (C - 1), 1]))
        e3d = np.arccos(e3d)
        scores.append(np.linalg.norm(e) / M + np.linalg.norm(e3d) / N)
        Rs.append(R); ts.append(t); sn.append(a); sm.append(b)

    # the reference's full loop; capture the inputs of its final inlier fit to recover the masks
    calls = []
    orig = rc.calib_linear

    def spy(v, nn):
        calls.append((v.shape[1], nn.shape[1], v.copy(), nn.copy()))
        return orig(v, nn)

    rc.calib_linear = spy
    try:
        with quiet(), contextlib.redirect_stderr(io.StringIO()):
            R_fin, t_fin, _ = rc.calib_linear_ransac(vc, n, sc.K, n_iter, th_2d, th_3d, seed)
    finally:
        rc.calib_linear = orig
    v_in, n_in = calls[-1][2], calls[-1][3]
    # inlier columns are an order-preserving subset: recover masks by exact column matching
    mask_3d = np.zeros(N, bool); j = 0
    for i in range(N):
        if j < v_in.shape[1] and np.array_equal(vc[:, i], v_in[:, j]):
            mask_3d[i] = True; j += 1
    mask_2d = np.zeros(M, bool); j = 0
    for i in range(M):
        if j < n_in.shape[1] and np.array_equal(n[:, i], n_in[:, j]):
            mask_2d[i] = True; j += 1
    assert mask_3d.sum() == v_in.shape[1] and mask_2d.sum() == n_in.shape[1]
    np.savez_compressed(os.path.join(OUT, f"ransac_{name}.npz"),
                        p3d=sc.p3d, s3d=sc.s3d, p2d=sc.p2d, s2d=sc.s2d, K=sc.K, vc=vc, n=n,
                        n_iter=n_iter, seed=seed, th_2d=th_2d, th_3d=th_3d,
                        scores=np.array(scores), R_hyp=np.array(Rs), t_hyp=np.array(ts),
                        sn=np.array(sn), sm=np.array(sm), best=int(np.argmin(scores)),
                        mask_2d=mask_2d, mask_3d=mask_3d, R_fin=R_fin, t_fin=t_fin,
                        versions=str(versions()))
    print("wrote ransac", name, "N_o =", N, "M =", M, "best =", int(np.argmin(scores)),
          "inliers", mask_2d.sum(), mask_3d.sum())


def gen_cli(name="cli_c4"):
    """File-level golden: a tiny scene written in the reference's directory layout by dataio.write_scene, read back
    by the reference's own util.load_eldersim and pushed through its main_ransac (calib_ransac.py:212-277)."""
    import tempfile
    dataio = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200.dataio")
    _, rc, ru = ref_shim.load()
    kw = dict(C=4, N=24, seed=31, p_drop_2d=0.1, p_drop_3d=0.05)
    sc = scene_mod.make_scene(**kw)
    aid, pid, gid, frame_skip, n_iter, seed = 23, 102, 3, 2, 5, 0
    with tempfile.TemporaryDirectory() as tmp:
        d = os.path.join(tmp, "noise_3_0")
        dataio.write_scene(d, sc, aid, pid, gid)
        loaded = ru.load_eldersim(d, gid, aid, pid)
        with quiet(), contextlib.redirect_stderr(io.StringIO()):
            R, t, _, _, camid, K = rc.main_ransac(d, gid, aid, pid, ru.OP_BONE, ru.OP_KEY_SUB, frame_skip=frame_skip,
                                                  n_iter=n_iter, th_2d=50.0, th_3d=5.0 / 180 * np.pi, seed=seed)
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), scene_kw=str(kw), aid=aid, pid=pid, gid=gid,
                        frame_skip=frame_skip, n_iter=n_iter, seed=seed, R_ransac=R, t_ransac=t, CAMID=camid, K=K,
                        p3d=loaded[5], s3d=loaded[6], p2d=loaded[7], s2d=loaded[8], frames=loaded[9],
                        versions=str(versions()))
    print("wrote", name)


def gen_save_mask(name="save_mask_c4"):
    """Golden of the observation-mask writer: the reference's own ba.save_mask (ba.py:235-266) on
    the cli_c4 files with a handful of gross 2-D outliers injected, cameras = the scene's perturbed start.  Stores the
    scores and poses the reference wrote to 2d_joint_mask/*.json (NaN = masked)."""
    import json
    import tempfile
    dataio = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200.dataio")
    rb, _, ru = ref_shim.load()
    kw = dict(C=4, N=24, seed=31, p_drop_2d=0.1, p_drop_3d=0.05)
    sc = scene_mod.make_scene(**kw)
    rng = np.random.default_rng(77)
    out_idx = np.stack([rng.integers(0, sc.C, 40), rng.integers(0, sc.N, 40), rng.integers(0, 25, 40)], axis=1)
    out_off = rng.normal(0, 40.0, (40, 2))            # px; some beyond, some within the 20 px threshold
    for (c, f, j), d in zip(out_idx, out_off):
        sc.p2d[c, f, j] += d
    aid, pid, gid, th_mask = 23, 102, 3, 20
    R, t = sc.R_init, sc.t_init.reshape(-1, 3, 1)
    scores, poses = [], []
    with tempfile.TemporaryDirectory() as tmp:
        d = os.path.join(tmp, "noise_3_0")
        dataio.write_scene(d, sc, aid, pid, gid)
        rb.PREFIX, rb.GID, rb.AID, rb.PID, rb.TH_MASK = d, gid, aid, pid, th_mask
        with quiet():
            rb.save_mask(sc.K, R, t, False, sc.width, sc.height)
        for cid in sc.camid:
            with open(os.path.join(d, "2d_joint_mask", f"A{aid:03d}_P{pid:03d}_G{gid:03d}_C{int(cid):03d}.json")) as fp:
                data = json.load(fp)["data"]
            scores.append([fr["skeleton"][0]["score"] for fr in data])
            poses.append([fr["skeleton"][0]["pose"] for fr in data])
            frames = [fr["frame_index"] for fr in data]
    scores, poses = np.array(scores, dtype=np.float64), np.array(poses, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), scene_kw=str(kw), out_idx=out_idx, out_off=out_off, aid=aid,
                        pid=pid, gid=gid, th_mask=th_mask, R=R, t=t, scores=scores, poses=poses, frames=np.array(frames),
                        versions=str(versions()))
    print("wrote", name, "masked", int(np.isnan(scores).sum()), "of", scores.size)


def gen_to_theta(name="to_theta_nonortho"):
    """ba.to_theta / from_theta (ba.py:18-31, cv2.Rodrigues) on NON-orthonormal rotations — what linear_*/ransac_*.json
    carry (calib_ransac.py:64-70): the reference's own theta and the R it unpacks again."""
    rb, _, ru = ref_shim.load()
    gl = np.load(os.path.join(OUT, "linear_c4.npz"))
    gr = np.load(os.path.join(OUT, "ransac_c4.npz"))
    Rs = np.concatenate([gl["R_linear"], gr["R_fin"], gr["R_hyp"][:4].reshape(-1, 3, 3)])
    ts = np.concatenate([gl["t_linear"].reshape(-1, 3, 1), gr["t_fin"].reshape(-1, 3, 1), gr["t_hyp"][:4].reshape(-1, 3, 1)])
    x = np.random.default_rng(5).normal(size=(7, 3))
    theta = rb.to_theta(Rs, ts, x)
    R2, t2, x2 = rb.from_theta(theta, len(Rs))
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), R=Rs, t=ts, x=x, theta=theta, R_back=R2, t_back=t2, x_back=x2,
                        versions=str(versions()))
    print("wrote", name, "max |RR^T - I| =", np.max(np.abs(Rs @ Rs.transpose(0, 2, 1) - np.eye(3))))


def gen_linear(name="linear_c4"):
    """File-level golden of the non-RANSAC linear calibration: the cli_c4 scene pushed through the reference's
    own calib_linear.main_linear (calib_linear.py:138-215)."""
    import tempfile
    dataio = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200.dataio")
    _, rc, ru = ref_shim.load()
    from importlib import util as _iu
    spec = _iu.spec_from_file_location("ref_calib_linear", os.path.join(ref_shim.REF, "calib_linear.py"))
    rl = _iu.module_from_spec(spec)            # the reference's file, not this repo's drop-in of the same name
    import types
    sys.modules["util"] = ru                              # its `from util import *` / `from argument import ...`
    pc = types.ModuleType("pycalib"); pc.calib = types.ModuleType("pycalib.calib")
    pc.calib.triangulate = ru.triangulate                 # vendored identical copy, util.py:240-258
    sys.modules["pycalib"], sys.modules["pycalib.calib"] = pc, pc.calib
    sys.path.insert(0, ref_shim.REF)
    argv, sys.argv = sys.argv, ["x"]
    try:
        spec.loader.exec_module(rl)
    finally:
        sys.argv = argv
        sys.path.remove(ref_shim.REF)
        for k in ("util", "argument", "pycalib", "pycalib.calib"):
            sys.modules.pop(k, None)
    kw = dict(C=4, N=24, seed=31, p_drop_2d=0.1, p_drop_3d=0.05)
    sc = scene_mod.make_scene(**kw)
    aid, pid, gid, frame_skip = 23, 102, 3, 2
    with tempfile.TemporaryDirectory() as tmp:
        d = os.path.join(tmp, "noise_3_0")
        dataio.write_scene(d, sc, aid, pid, gid)
        loaded = ru.load_eldersim(d, gid, aid, pid)
        with quiet(), contextlib.redirect_stderr(io.StringIO()):
            R, t, X, e, camid, K = rl.main_linear(d, gid, aid, pid, ru.OP_BONE, ru.OP_KEY_SUB, False,
                                                  frame_skip=frame_skip)
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), scene_kw=str(kw), aid=aid, pid=pid, gid=gid,
                        frame_skip=frame_skip, R_linear=R, t_linear=t, X_linear=X, e_linear=e, CAMID=camid, K=K,
                        p3d=loaded[5], s3d=loaded[6], p2d=loaded[7], s2d=loaded[8], frames=loaded[9],
                        versions=str(versions()))
    print("wrote", name)


def gen_eval(name="eval_c4", cams="init"):
    """Golden of the evaluation metrics: the reference's own eval.align / eval.calc_repro_error / util.eval_R /
    util.eval_t / util.triangulate_with_conf, strung together as eval.eval_main does (eval.py:126-150), plus the
    JSON summary eval_main writes.  pycalib.calib.absolute_orientation is not installable here: it is stubbed with
    the restatement in calib_b200/eval.py, so THAT function is not pinned by this golden (everything around it is)."""
    import tempfile
    import types
    from importlib import util as _iu
    my_eval = importlib.import_module("extrinsic-camera-calibration-from-a-moving-person_b200.eval")
    _, rc, ru = ref_shim.load()
    stubs = {}
    def stub(nm, **attrs):
        m = types.ModuleType(nm); m.__dict__.update(attrs); stubs[nm] = m; sys.modules[nm] = m; return m
    stub("matplotlib"); stub("matplotlib.pyplot"); stub("mpl_toolkits"); stub("mpl_toolkits.mplot3d", Axes3D=object)
    pc = stub("pycalib"); pc.calib = stub("pycalib.calib", absolute_orientation=my_eval.absolute_orientation,
                                          triangulate=ru.triangulate)
    pc.plot = stub("pycalib.plot", plotCamera=lambda *a, **k: None)
    sys.modules["util"] = ru
    sys.path.insert(0, ref_shim.REF)
    argv, sys.argv = sys.argv, ["x"]
    try:
        spec = _iu.spec_from_file_location("ref_eval", os.path.join(ref_shim.REF, "eval.py"))
        re_ = _iu.module_from_spec(spec)
        spec.loader.exec_module(re_)
    finally:
        sys.argv = argv
        sys.path.remove(ref_shim.REF)
        for k in list(stubs) + ["util", "argument"]:
            sys.modules.pop(k, None)
    kw = dict(C=4, N=24, seed=31, p_drop_2d=0.1, p_drop_3d=0.05)
    sc = scene_mod.make_scene(**kw)
    frame_skip, scale, width, height = 2, 0.01, 640, 360
    gp2d, gs2d = np.array(sc.p2d), np.array(sc.s2d)
    gp3d_w = np.array(sc.X_w)
    gt = sc.t_gt.reshape(-1, 3, 1)
    R, t = sc.R_init, sc.t_init.reshape(-1, 3, 1)
    if cams == "linear":       # non-orthonormal R, as run_all.sh feeds eval.py with linear_*/ransac_*.json
        gl = np.load(os.path.join(OUT, "linear_c4.npz"))
        R, t = gl["R_linear"], gl["t_linear"].reshape(-1, 3, 1)
        # the linear model lives in camera 0's frame at unit baseline: evaluate it as eval.py would, as is
    with quiet():
        x_all = ru.triangulate_with_conf(gp2d, gs2d, sc.K, R, t, (gs2d > 0))
        N, J = x_all.shape[:2]
        nR, nt, _, _, np3d_w = re_.align(sc.R_gt, gt, R, t, x_all.reshape(-1, 3))
        np3d_w = np3d_w.reshape(N, J, 3)
        mask_3dw = np.all(~np.isnan(gp3d_w), axis=2)
        E_x = np.linalg.norm(np3d_w[mask_3dw] - gp3d_w[mask_3dw], axis=1) * scale
        E_R = np.array([ru.eval_R(sc.R_gt[i], nR[i]) for i in range(len(R))])
        E_t = np.array([ru.eval_t(gt[i], nt[i]) for i in range(len(R))]) * scale
        E_p, _, _, _ = re_.calc_repro_error(gs2d > 0, frame_skip, x_all, gp2d, sc.K, height, width, R, t)
        with tempfile.TemporaryDirectory() as tmp:
            jo = os.path.join(tmp, "e.json")
            re_.eval_main(sc.K, R, t, sc.R_gt, gt, gp2d, gs2d, gp3d_w, frame_skip, width, height, jo, scale)
            summary = open(jo).read()
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), scene_kw=str(kw), frame_skip=frame_skip, scale=scale,
                        E_R=E_R, E_t=E_t, E_p=E_p, E_x=E_x, nR=nR, nt=nt, summary=summary, R=R, t=t, versions=str(versions()))
    print("wrote", name)


def main():
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 2 and sys.argv[1] == "tight":
        kw, nj, l1, l2 = TIGHT_CASES[sys.argv[2]]
        big = kw["N"] >= 40
        gen_ba_tight(sys.argv[2], kw, list(range(nj)), l1, l2, max_nfev=150 if big else 300,
                     jacs=("3-point",) if big else ("2-point", "3-point"))
        return
    if len(sys.argv) > 1 and sys.argv[1] == "ransac_long":
        # long enough to meet ill-conditioned minimal samples (the reference never rejects them,
        # calib_ransac.py:99-104): exercises the deferred Jacobi fall-back of the CUDA fit
        gen_ransac("c4_long", dict(C=4, N=40, seed=21, p_drop_2d=0.1, p_drop_3d=0.05), 1500, 0, 50.0, 5.0)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "to_theta":
        gen_to_theta()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "save_mask":
        gen_save_mask()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "eval":
        gen_eval()
        gen_eval("eval_c4_linear", "linear")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "linear":
        gen_linear()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "cli":
        gen_cli()
        return
    gen_ba("small", dict(C=4, N=6, seed=11), list(range(17)), 1.0, 1.0, run_main=True)
    gen_ba("partial", dict(C=5, N=8, seed=12, score_mode="uniform", p_drop_2d=0.3, p_drop_3d=0.1),
           list(range(17)), 3.0, 7.0)
    synadl_joints = list(range(23)) + [22, 23, 24]          # config.yaml:10-36 lists 22 twice
    gen_ba("synadl26", dict(C=4, N=5, seed=13), synadl_joints, 1.0, 1.0)
    gen_ba("gafa_lambda", dict(C=3, N=7, seed=14, score_mode="uniform"), list(range(17)), 1.0, 1e5)
    gen_ransac("c4", dict(C=4, N=40, seed=21, p_drop_2d=0.1, p_drop_3d=0.05), 12, 0, 50.0, 5.0)
    gen_ransac("c3", dict(C=3, N=30, seed=22, sigma_px=1.0), 8, 7, 50.0, 5.0)
    gen_cli()
    gen_linear()
    gen_eval()
    gen_eval('eval_c4_linear', 'linear')
    gen_save_mask()
    gen_to_theta()
    for nm, (kw, nj, l1, l2) in TIGHT_CASES.items():
        big = kw['N'] >= 40
        gen_ba_tight(nm, kw, list(range(nj)), l1, l2, max_nfev=150 if big else 300,
                     jacs=('3-point',) if big else ('2-point', '3-point'))
    gen_ransac('c4_long', dict(C=4, N=40, seed=21, p_drop_2d=0.1, p_drop_3d=0.05), 1500, 0, 50.0, 5.0)


if __name__ == "__main__":
    main()
