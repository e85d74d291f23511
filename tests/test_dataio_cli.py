"""File formats and command-line surface either side of the hot path.  CPU part: dataio round trips against what
the reference's own util.load_eldersim returned for the same files (tests/golden/cli_c4.npz).  GPU part: the
calib_ransac.py -> ba.py chain through the drop-in scripts at the repo root."""
import ast
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import calib_b200  # noqa: F401
from calib_b200 import dataio, scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _golden_scene(golden_dir):
    g = np.load(os.path.join(golden_dir, "cli_c4.npz"))
    sc = scene.make_scene(**ast.literal_eval(str(g["scene_kw"])))
    return g, sc


def test_scene_files_load_like_the_reference(golden_dir, tmp_path):
    g, sc = _golden_scene(golden_dir)
    d = str(tmp_path / "noise_3_0")
    dataio.write_scene(d, sc, int(g["aid"]), int(g["pid"]), int(g["gid"]))
    camid, K, R, t, p3d_w, p3d, s3d, p2d, s2d, frames = dataio.load_scene(d, int(g["gid"]), int(g["aid"]), int(g["pid"]))
    assert np.array_equal(camid, g["CAMID"]) and np.array_equal(K, g["K"])
    for mine, ref in ((p3d, g["p3d"]), (s3d, g["s3d"]), (p2d, g["p2d"]), (s2d, g["s2d"]), (frames, g["frames"])):
        assert mine.shape == ref.shape and np.array_equal(mine, ref)
    assert np.array_equal(R, sc.R_gt) and t.shape == (4, 3, 1) and np.array_equal(p3d_w, sc.X_w)
    # NumPy side channel gives the same arrays
    d2 = str(tmp_path / "npz" / "noise_3_0")
    dataio.write_scene(d2, sc, 23, 102, 3, npz=True)
    again = dataio.load_scene(d2, 3, 23, 102)
    assert np.array_equal(again[7], p2d) and np.array_equal(again[5], p3d) and np.array_equal(again[9], frames)


def test_pose_files_keep_nan_scores_and_frame_subsets(tmp_path):
    p = np.arange(2 * 25 * 2, dtype=np.float64).reshape(2, 25, 2)
    s = np.ones((2, 25)); s[1, 3] = np.nan
    f = str(tmp_path / "x.json")
    dataio.save_poses(f, [4, 9], p, s)
    fr, pose, score = dataio.load_poses(f)
    assert list(fr) == [4, 9] and np.array_equal(pose.reshape(2, 25, 2), p)
    assert np.isnan(score[1, 3]) and np.array_equal(np.isnan(score), np.isnan(s))
    raw = json.load(open(f))
    assert set(raw["data"][0]) == {"frame_index", "skeleton"} and set(raw["data"][0]["skeleton"][0]) == {"pose", "score"}


def test_camera_file_keys_and_shapes(tmp_path):
    f = str(tmp_path / "c.json")
    R = np.stack([np.eye(3)] * 3); t = np.arange(9.0).reshape(3, 3, 1)
    dataio.save_cameras(f, np.array([1, 2, 5]), R * 2, R, t)
    raw = json.load(open(f))
    assert list(raw) == ["CAMID", "K", "R_w2c", "t_w2c"]
    camid, K, R2, t2 = dataio.load_cameras(f)
    assert t2.shape == (3, 3, 1) and np.array_equal(t2, t) and np.array_equal(camid, [1, 2, 5])


def test_cli_flags_follow_the_reference_parser():
    from calib_b200 import cli
    a = cli.parse_args([])
    assert (a.dataset, a.aid, a.pid, a.gid, a.frame_skip, a.ransac_niter, a.ransac_seed, a.target) == \
        ("SynADL", 23, 102, 3, 15, 100, 0, "noise_3_0")
    assert a.ba_lambda1 == 1.0 and a.ba_lambda2 == 1.0 and a.th_obs_mask == 20 and not a.obs_mask
    a = cli.parse_args("--obs_mask true --save_obs_mask False --ba_lambda2 1e5 --calib 1 2".split())
    assert a.obs_mask == 1 and a.save_obs_mask == 0 and a.ba_lambda2 == 1e5 and a.calib == [1, 2]
    cfg = cli.load_config("SynADL", os.path.join(ROOT, "config", "config.yaml"))
    assert cfg["available_joints"].count(22) == 2 and len(cfg["available_joints"]) == 26
    assert cfg["ransac_th_2d"] == 50.0 and cfg["ransac_th_3d"] == 5.0 and (cfg["width"], cfg["height"]) == (640, 360)


@pytest.mark.gpu
def test_cli_chain_ransac_then_ba(golden_dir, tmp_path):
    g, sc = _golden_scene(golden_dir)
    prefix = str(tmp_path / "A023_P102_G003")
    dataio.write_scene(os.path.join(prefix, "noise_3_0"), sc, 23, 102, 3)
    env = dict(os.environ, RANSAC_NITER=str(int(g["n_iter"])), PYTHONPATH=ROOT)
    out = subprocess.run(["bash", os.path.join(ROOT, "calib_ransac.sh"), prefix, "23", "102", "3", "noise_3_0",
                          str(int(g["frame_skip"])), "SynADL"], cwd=ROOT, env=env, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "dataset=SynADL" in out.stdout and "vc=(4," in out.stdout and "rejected, th=" in out.stdout
    camid, K, R, t = dataio.load_cameras(os.path.join(prefix, "results", "ransac_3_0.json"))
    # same files through the reference's own main_ransac (oracle/gen_golden.py gen_cli)
    assert np.array_equal(camid, g["CAMID"])
    assert np.max(np.abs(R - g["R_ransac"])) < 1e-9 and np.max(np.abs(t - g["t_ransac"])) < 1e-9
    out = subprocess.run(["bash", os.path.join(ROOT, "ba.sh"), prefix, "23", "102", "3", str(int(g["frame_skip"])), "1.", "1.",
                          "ransac_3_0", "SynADL", "false", "true"], cwd=ROOT, env=env, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "target BA=" in out.stdout and "Initial mean NLL:" in out.stdout and "save obs. mask" in out.stdout
    camid, K, R2, t2 = dataio.load_cameras(os.path.join(prefix, "results", "ransac_3_0_ba.json"))
    assert R2.shape == (4, 3, 3) and t2.shape == (4, 3, 1)
    assert np.max(np.abs(R2 @ np.transpose(R2, (0, 2, 1)) - np.eye(3))) < 1e-12          # BA returns true rotations
    # BA moved the cameras towards the ground truth relative pose (sanity)
    rel = lambda Rs: Rs[1] @ Rs[0].T
    err0 = np.linalg.norm(rel(R) - rel(sc.R_gt)); err1 = np.linalg.norm(rel(R2) - rel(sc.R_gt))
    assert err1 < err0
    masked = os.path.join(prefix, "noise_3_0", "2d_joint_mask", dataio.joint_file(23, 102, 3, 21))
    fr, pose, score = dataio.load_poses(masked)
    assert len(fr) == sc.N and pose.shape == (sc.N, 50) and score.shape == (sc.N, 25)


@pytest.mark.gpu
def test_calib_linear_dropin_vs_reference_golden(golden_dir, tmp_path):
    """calib_linear.sh / calib_linear.main_linear on the files the reference's own main_linear read
    (oracle/gen_golden.py gen_linear): cameras 1e-9, re-triangulated points 1e-8 relative, its reprojection figure."""
    g = np.load(os.path.join(golden_dir, "linear_c4.npz"))
    sc = scene.make_scene(**ast.literal_eval(str(g["scene_kw"])))
    prefix = str(tmp_path / "A023_P102_G003")
    dataio.write_scene(os.path.join(prefix, "noise_3_0"), sc, 23, 102, 3)
    from calib_b200 import calib_linear as cl
    from calib_b200.ba import OP_BONE
    R, t, X, e, camid, K = cl.main_linear(os.path.join(prefix, "noise_3_0"), 3, 23, 102, OP_BONE, None, False,
                                          frame_skip=int(g["frame_skip"]))
    assert np.array_equal(camid, g["CAMID"]) and np.array_equal(K, g["K"])
    assert np.max(np.abs(R - g["R_linear"])) < 1e-9 and np.max(np.abs(t - g["t_linear"])) < 1e-9
    assert X.shape == g["X_linear"].shape
    assert np.max(np.abs(X - g["X_linear"])) < 1e-8 * np.max(np.abs(g["X_linear"]))
    assert abs(e - float(g["e_linear"])) < 1e-8 * float(g["e_linear"])
    env = dict(os.environ, PYTHONPATH=ROOT)
    out = subprocess.run(["bash", os.path.join(ROOT, "calib_linear.sh"), prefix, "23", "102", "3", "noise_3_0",
                          str(int(g["frame_skip"])), "SynADL", "false"], cwd=ROOT, env=env, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "dataset=SynADL" in out.stdout and "vc=(4," in out.stdout
    camid2, K2, R2, t2 = dataio.load_cameras(os.path.join(prefix, "results", "linear_3_0.json"))
    assert np.array_equal(camid2, g["CAMID"]) and np.max(np.abs(R2 - g["R_linear"])) < 1e-9
    assert np.max(np.abs(t2 - g["t_linear"])) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["eval_c4", "eval_c4_linear"])
def test_eval_metrics_vs_reference_golden(golden_dir, tmp_path, capsys, case):
    """eval.eval_main on the arrays the reference's own eval.align / calc_repro_error / util.eval_R / eval_t /
    triangulate_with_conf were run on (oracle/gen_golden.py gen_eval; pycalib.absolute_orientation restated there
    and here alike — that one function is unpinned): error vectors 1e-8 relative, identical JSON summary."""
    g = np.load(os.path.join(golden_dir, f"{case}.npz"))     # _linear: non-orthonormal R (linear_*.json), util.invRT_batch
    sc = scene.make_scene(**ast.literal_eval(str(g["scene_kw"])))
    from calib_b200 import eval as ev
    jo = str(tmp_path / "e.json")
    E_R, E_t, E_p, E_x = ev.eval_main(sc.K, g["R"], g["t"], sc.R_gt, sc.t_gt.reshape(-1, 3, 1),
                                      np.array(sc.p2d), np.array(sc.s2d), np.array(sc.X_w), int(g["frame_skip"]), 640, 360,
                                      jo, float(g["scale"]))
    for mine, ref in ((E_R, g["E_R"]), (E_t, g["E_t"]), (E_p, g["E_p"]), (E_x, g["E_x"])):
        assert mine.shape == ref.shape
        assert np.max(np.abs(mine - ref)) < 1e-8 * max(1.0, np.nanmax(np.abs(ref)))
    assert json.load(open(jo)) == json.loads(str(g["summary"]))
    out = capsys.readouterr().out
    assert "       Min  Median  Mean  Max" in out and "E_R:" in out and "E_x:" in out
    from calib_b200 import ba
    prob = ba.BAProblem(sc.K, np.array(sc.p2d), np.array(sc.s2d), None, None, ba.OP_BONE, 0.0, 0.0)
    x_all = prob.triangulate(g["R"], g["t"])
    prob.close()
    E_p2, x_gt_all, _, _ = ev.calc_repro_error(np.array(sc.s2d) > 0, int(g["frame_skip"]), x_all, np.array(sc.p2d), sc.K,
                                               360, 640, g["R"], g["t"])
    assert np.allclose(E_p2, g["E_p"], rtol=1e-8, atol=1e-10) and x_gt_all.shape == np.array(sc.p2d).shape


@pytest.mark.gpu
def test_save_mask_vs_reference_golden(golden_dir, tmp_path):
    """cli.save_mask against the files the reference's own ba.save_mask (ba.py:235-266) wrote for the same scene,
    cameras and threshold (oracle/gen_golden.py gen_save_mask): which scores became NaN is bit-equal, every other
    score and every pose value is identical, frame indices too."""
    from calib_b200 import cli
    g = np.load(os.path.join(golden_dir, "save_mask_c4.npz"))
    sc = scene.make_scene(**ast.literal_eval(str(g["scene_kw"])))
    for (c, f, j), d in zip(g["out_idx"], g["out_off"]):
        sc.p2d[c, f, j] += d
    aid, pid, gid = int(g["aid"]), int(g["pid"]), int(g["gid"])
    d = str(tmp_path / "noise_3_0")
    dataio.write_scene(d, sc, aid, pid, gid)
    n_far = cli.save_mask(d, gid, aid, pid, sc.K, g["R"], g["t"], False, int(g["th_mask"]), sc.width, sc.height)
    scores, poses = [], []
    for cid in sc.camid:
        fr, pose, score = dataio.load_poses(os.path.join(d, "2d_joint_mask", dataio.joint_file(aid, pid, gid, int(cid))))
        assert np.array_equal(fr, g["frames"])
        scores.append(score); poses.append(pose)
    scores, poses = np.array(scores), np.array(poses)
    assert scores.shape == g["scores"].shape and poses.shape == g["poses"].shape
    assert np.array_equal(np.isnan(scores), np.isnan(g["scores"]))
    assert int(np.isnan(g["scores"]).sum()) > 10 and n_far >= int(np.isnan(g["scores"]).sum()) - int(np.isnan(np.array(sc.s2d)).sum())
    assert np.array_equal(np.nan_to_num(scores, nan=-1.0), np.nan_to_num(g["scores"], nan=-1.0))
    assert np.array_equal(poses, g["poses"])
