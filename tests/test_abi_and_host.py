"""CPU-only checks: the C-ABI library loads and exports every symbol declared in include/*.h; the NumPy-compatible
sample stream (host C++); and the device numerics of the RANSAC path run through the host emulation build
(tests/hostemu, same __host__ __device__ code as the kernels) against golden vectors of the unmodified reference."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import calib_b200  # noqa: F401
from calib_b200 import _lib, _rsx_sigs
from oracle import ransac_oracle as ro

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = C.POINTER(C.c_double)


def dp(a):
    return a.ctypes.data_as(P)


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b((?:bax|rsx)_[a-z0-9_]+)\s*\(", txt)))


@pytest.mark.parametrize("header", ["bax.h", "rsx.h"])
def test_library_exports_every_declared_symbol(header):
    L = _lib.lib()
    names = _declared(header)
    assert len(names) >= 8
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/{header} but not exported by libbax.so"
    bound = set(_lib.exported_symbols()) | set(_rsx_sigs.exported_symbols())
    assert set(names) <= bound, f"ctypes binding missing for {set(names) - bound}"


def test_new_entry_points_reject_bad_arguments_without_touching_the_gpu():
    """Argument checks of the round-2 entry points (no device work happens before them): error code + message."""
    import ctypes as C_
    L = _lib.lib()
    assert L.bax_batch_solve(None) == -1 and b"bax_batch_solve" in L.bax_last_error()
    a = _lib.BatchArgs()
    a.n_problems, a.C, a.N, a.J = 3, 4, 10, 17                    # no pointers filled in
    assert L.bax_batch_solve(C_.byref(a)) == -1
    st = (C_.c_int64 * 4)()
    assert L.rsx_fit_stats(None, st, None) == -1 and b"rsx_fit_stats" in L.bax_last_error()
    assert L.bax_comm_stats(None, st) == -1
    assert L.bax_allreduce_attach(None, None) == -1
    assert L.bax_set_blocking_sync(None, 1) == -1
    assert L.bax_comm_unique_id(None) != 0
    w = (C_.c_double * 6)()
    assert L.rsx_calib_linear_eigs(None, w) == -1
    assert L.bax_batch_release() == 0                            # nothing parked: a no-op


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from calib_b200 import ba, calib_ransac
    with pytest.raises(_lib.BaxError):
        ba.BAProblem(np.eye(3)[None], np.zeros((1, 1, 1, 2)), np.ones((1, 1, 1)), None, None)
    with pytest.raises(_lib.BaxError):
        calib_ransac.RansacProblem(np.zeros((2, 10, 3)), np.zeros((2, 10, 3)))


@pytest.mark.parametrize("N,M,H,seed", [(360, 750, 400, 0), (12000, 15000, 300, 7), (10, 10, 50, 1), (11, 4000, 64, 123)])
def test_sample_stream_matches_numpy(N, M, H, seed):
    from calib_b200 import calib_ransac as cr
    sn, sm = cr.sample_indices(N, M, H, seed)
    rn, rm = ro.sample_indices(N, M, H, seed)          # np.random.default_rng(seed).choice(...), calib_ransac.py:148-151
    assert np.array_equal(sn, rn) and np.array_equal(sm, rm)


@pytest.mark.parametrize("case", ["c4", "c3"])
def test_sample_stream_matches_reference_golden(golden_dir, case):
    from calib_b200 import calib_ransac as cr
    g = np.load(os.path.join(golden_dir, f"ransac_{case}.npz"))
    sn, sm = cr.sample_indices(g["vc"].shape[1], g["n"].shape[1], int(g["n_iter"]), int(g["seed"]))
    assert np.array_equal(sn, g["sn"]) and np.array_equal(sm, g["sm"])


def test_ransac_inputs_match_reference_golden(golden_dir):
    from calib_b200 import calib_ransac as cr
    from calib_b200.ba import OP_BONE
    for case in ("c4", "c3"):
        g = np.load(os.path.join(golden_dir, f"ransac_{case}.npz"))
        v, n = cr.ransac_inputs(g["p3d"], g["s3d"], g["p2d"], g["s2d"], g["K"], OP_BONE)
        assert np.array_equal(v, g["vc"]) and np.array_equal(n, g["n"])


def test_shard_ranges_cover_all_hypotheses():
    from calib_b200 import calib_ransac as cr
    for n_iter in (1, 7, 100000):
        for world in (1, 2, 3, 8):
            r = [cr.shard_range(n_iter, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n_iter
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))


def test_chunk_plan_covers_every_hypothesis_once_with_equal_shares():
    from calib_b200 import calib_ransac as cr
    for n_iter in (1, 7, 4097, 50000, 100000):
        for world in (1, 2, 3, 8):
            plan = cr.chunk_plan(n_iter, world)
            assert plan[0][0] == 0 and plan[-1][1] == n_iter
            assert all(plan[i][1] == plan[i + 1][0] for i in range(len(plan) - 1))
            assert all(0 <= o < world and b > a for a, b, o in plan)
            share = [sum(b - a for a, b, o in plan if o == r) for r in range(world)]
            assert sum(share) == n_iter
            if n_iter >= 50000:
                assert max(share) - min(share) <= len(plan)          # equal up to the rounding of the chunk size


def test_eval_alignment_recovers_a_similarity_transform():
    """eval.align / absolute_orientation (eval.py:29-63; pycalib restated, parity unpinned): cameras and points moved
    by a known similarity transform come back exactly; E_R / E_t are then zero."""
    from calib_b200 import eval as ev
    from calib_b200 import scene
    sc = scene.make_scene(C=5, N=4, seed=3)
    rng = np.random.default_rng(0)
    A = scene.rodrigues(rng.normal(0, 1.0, 3)); s0 = 2.7; d = rng.normal(0, 1, 3)
    # estimated frame: X_e = (A^T (X_g - d)) / s0  ->  R_e = R_g A, t_e = (t_g + R_g d) / s0
    R_e = np.array([R @ A for R in sc.R_gt])
    t_e = np.array([(t.reshape(3) + R @ d) / s0 for R, t in zip(sc.R_gt, sc.t_gt)]).reshape(-1, 3, 1)
    X_g = rng.normal(0, 1, (30, 3)); X_e = (X_g - d) @ A / s0
    Rn, tn, _, _, Xn = ev.align(sc.R_gt, sc.t_gt.reshape(-1, 3, 1), R_e, t_e, X_e)
    assert np.max(np.abs(Rn - sc.R_gt)) < 1e-12 and np.max(np.abs(tn - sc.t_gt.reshape(-1, 3, 1))) < 1e-11
    assert np.max(np.abs(Xn - X_g)) < 1e-11
    assert max(ev.eval_R(a, b) for a, b in zip(sc.R_gt, Rn)) < 1e-7 and ev.eval_t(sc.t_gt[0], tn[0]) < 1e-11
    R, t, s = ev.absolute_orientation(X_e.T, X_g.T)
    assert abs(s - s0) < 1e-12 and np.max(np.abs(R - A)) < 1e-12 and np.max(np.abs(t - d)) < 1e-12


# ---------------------------------------------------------------- device math through the host emulation
@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "tests", "hostemu", "librsx_hostemu.so")
    src = os.path.join(ROOT, "tests", "hostemu", "rsx_hostemu.cu")
    if not os.path.isfile(so):
        subprocess.check_call(["bash", os.path.join(ROOT, "tests", "hostemu", "build.sh")])
    assert os.path.isfile(so) and os.path.isfile(src)
    return C.CDLL(so)


@pytest.mark.parametrize("n", [4, 10, 41, 42])
def test_emulated_cta_jacobi(emu, n):
    rng = np.random.default_rng(n)
    B = rng.normal(size=(n + 5 if n != 42 else 30, n))
    A0 = B.T @ B
    A, V = A0.copy(), np.zeros((n, n))
    sweeps = emu.emu_jacobi(n, dp(A), dp(V))
    assert sweeps < 40
    w = np.diag(A).copy()
    assert np.linalg.norm(V.T @ V - np.eye(n)) < 1e-12
    assert np.linalg.norm(A0 @ V - V * w) <= 1e-13 * np.linalg.norm(A0)
    assert np.max(np.abs(np.sort(w) - np.linalg.eigvalsh(A0))) <= 1e-13 * np.max(w)


def test_emulated_eig4(emu):
    rng = np.random.default_rng(0)
    for _ in range(200):
        B = rng.normal(size=(rng.integers(3, 9), 4))
        A = B.T @ B
        x = np.zeros(4)
        emu.emu_eig4(dp(np.ascontiguousarray(A)), dp(x))
        w, v = np.linalg.eigh(A)
        assert abs(abs(x @ v[:, 0]) - 1) < 1e-9 or w[1] - w[0] < 1e-6 * w[3]
        assert abs(x @ A @ x - w[0]) <= 1e-12 * w[3]


def test_emulated_eig4_inverse_iteration(emu):
    """Cholesky inverse iteration (the scoring kernel's DLT solver): random matrices (poor gaps -> Jacobi fall-back),
    DLT-like matrices with a small smallest eigenvalue, rank-deficient ones (3 rows) and NaN input."""
    rng = np.random.default_rng(1)
    for k in range(400):
        B = rng.normal(size=(rng.integers(3, 9), 4))
        if k % 2:                                   # triangulation-like: one direction nearly in the null space
            u, sv, vt = np.linalg.svd(rng.normal(size=(8, 4)), full_matrices=False)
            sv[3] = sv[2] * 10.0 ** rng.uniform(-8, -1)
            B = (u * sv) @ vt
        A = B.T @ B
        x = np.zeros(4)
        emu.emu_eig4_fast(dp(np.ascontiguousarray(A)), dp(x))
        w, v = np.linalg.eigh(A)
        assert abs(np.linalg.norm(x) - 1) < 1e-12
        assert abs(abs(x @ v[:, 0]) - 1) < 1e-9 or w[1] - w[0] < 1e-6 * w[3]
        assert abs(x @ A @ x - w[0]) <= 1e-12 * w[3]
        if k % 2:
            assert np.linalg.norm(x * np.sign(x @ v[:, 0]) - v[:, 0]) < 1e-12 * max(1.0, w[3] / (w[1] - w[0]))
    x = np.zeros(4)
    emu.emu_eig4_fast(dp(np.full((4, 4), np.nan)), dp(x))      # must terminate


@pytest.mark.parametrize("case", ["c4", "c3"])
def test_emulated_fit_score_masks_vs_reference_golden(emu, golden_dir, case):
    g = np.load(os.path.join(golden_dir, f"ransac_{case}.npz"))
    v, n = np.ascontiguousarray(g["vc"]), np.ascontiguousarray(g["n"])
    Cn, N, M = v.shape[0], v.shape[1], n.shape[1]
    pi = C.POINTER(C.c_int)
    models = []
    for h in range(int(g["n_iter"])):
        sv = np.ascontiguousarray(g["sn"][h], dtype=np.int32)
        sm = np.ascontiguousarray(g["sm"][h], dtype=np.int32)
        Rt = np.zeros((Cn, 3, 4))
        emu.emu_fit(Cn, C.c_longlong(N), C.c_longlong(M), dp(v), dp(n), sv.ctypes.data_as(pi), sm.ctypes.data_as(pi), dp(Rt))
        assert np.max(np.abs(Rt[:, :, :3] - g["R_hyp"][h])) < 1e-9
        assert np.max(np.abs(Rt[:, :, 3:] - g["t_hyp"][h])) < 1e-9
        parts = np.zeros(2)
        emu.emu_score(Cn, C.c_longlong(N), C.c_longlong(M), dp(v), dp(n), dp(Rt), dp(parts), None, None, None,
                      C.c_double(0), C.c_double(0), None, None)
        score = np.sqrt(parts[0]) / M + np.sqrt(parts[1]) / N
        assert abs(score - g["scores"][h]) <= 1e-9 * g["scores"][h]
        models.append((score, Rt))
    best = int(np.argmin([m[0] for m in models]))
    assert best == int(g["best"])
    m2, m3 = np.zeros(M, np.uint8), np.zeros(N, np.uint8)
    K00 = np.ascontiguousarray(g["K"][:, 0, 0])
    pu = C.POINTER(C.c_ubyte)
    parts = np.zeros(2)
    emu.emu_score(Cn, C.c_longlong(N), C.c_longlong(M), dp(v), dp(n), dp(models[best][1]), dp(parts), None, None, dp(K00),
                  C.c_double(float(g["th_2d"])), C.c_double(float(g["th_3d"])), m2.ctypes.data_as(pu), m3.ctypes.data_as(pu))
    assert np.array_equal(m2.astype(bool), g["mask_2d"]) and np.array_equal(m3.astype(bool), g["mask_3d"])


def test_package_to_theta_vs_reference_on_non_orthonormal_rotations(golden_dir):
    """calib_b200.ba.to_theta / from_theta (the PACKAGE's host functions, not the oracle's) against the reference's
    ba.to_theta / from_theta (ba.py:18-31) on the non-orthonormal R of linear_*/ransac_* models: cv2.Rodrigues
    projects them on SO(3) by SVD first (oracle/gen_golden.py gen_to_theta)."""
    from calib_b200 import ba
    g = np.load(os.path.join(golden_dir, "to_theta_nonortho.npz"))
    assert np.max(np.abs(g["R"] @ g["R"].transpose(0, 2, 1) - np.eye(3))) > 1e-3      # really non-orthonormal
    th = ba.to_theta(g["R"], g["t"], g["x"])
    assert th.shape == g["theta"].shape and np.max(np.abs(th - g["theta"])) < 1e-12
    R, t, x = ba.from_theta(g["theta"], len(g["R"]))
    assert np.max(np.abs(R - g["R_back"])) < 1e-12 and np.array_equal(t, g["t_back"]) and np.array_equal(x, g["x_back"])


@pytest.mark.parametrize("case", ["eval_c4", "eval_c4_linear"])
def test_eval_align_vs_reference_golden(golden_dir, case):
    """eval.align (eval.py:29-63) around the (unpinned) absolute_orientation: aligned cameras equal to the reference's,
    including util.invRT_batch's 4x4 inverse for the non-orthonormal R of linear_*.json."""
    import ast
    from calib_b200 import eval as ev, scene
    g = np.load(os.path.join(golden_dir, f"{case}.npz"))
    sc = scene.make_scene(**ast.literal_eval(str(g["scene_kw"])))
    nR, nt, _, _, _ = ev.align(sc.R_gt, sc.t_gt.reshape(-1, 3, 1), g["R"], g["t"], np.zeros((1, 3)))
    assert np.max(np.abs(nR - g["nR"])) < 1e-12 and np.max(np.abs(nt - g["nt"])) < 1e-12


def test_sample_plan_caches_and_extends_the_reference_stream():
    """SamplePlan (calib_ransac.py host side): the cached, pinned-when-possible copy of the reference's sample stream
    (calib_ransac.py:148-151) — generated lazily in chunks, extended on demand, handed back on the next call with the same
    (seed, N, M), and always equal to NumPy's own stream."""
    from calib_b200 import calib_ransac as cr
    cr.SamplePlan._cache.clear()
    N, M, seed = 149, 325, 5
    rng = np.random.default_rng(seed)
    ref_n, ref_m = [], []
    for _ in range(300):
        ref_n.append(rng.choice(N, size=10, replace=False)); ref_m.append(rng.choice(M, size=10, replace=False))
    ref_n, ref_m = np.array(ref_n), np.array(ref_m)
    plan = cr.SamplePlan.get(N, M, seed, 100)
    assert plan.done == 0
    plan.ensure(40); plan.ensure(100)                      # two chunks
    assert plan.done == 100 and np.array_equal(plan.sn[:100], ref_n[:100]) and np.array_equal(plan.sm[:100], ref_m[:100])
    again = cr.SamplePlan.get(N, M, seed, 300)             # same key: the same plan, grown, earlier samples kept
    assert again is plan and plan.done == 100
    assert plan.ensure(300) > 0.0 and plan.ensure(300) == 0.0
    assert np.array_equal(plan.sn[:300], ref_n) and np.array_equal(plan.sm[:300], ref_m)
    other = cr.SamplePlan.get(N, M, seed + 1, 10)
    assert other is not plan
    for k in range(6):                                     # bounded cache
        cr.SamplePlan.get(N + k + 1, M, seed, 10)
    assert len(cr.SamplePlan._cache) <= cr.SamplePlan.MAX_CACHED
    sn, sm = cr.sample_indices(N, M, 300, seed)
    assert np.array_equal(sn, ref_n) and np.array_equal(sm, ref_m)
