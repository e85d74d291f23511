"""world_size-2 `gloo` tests (CPU) of the host-side multi-GPU logic: hypothesis sharding + winner agreement of the
RANSAC path, and the frame partition / packed-buffer summation identity the BA path relies on."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import calib_b200  # noqa: F401


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    mgr = mp.Manager(); ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def _winner_case(rank, world):
    from calib_b200 import calib_ransac as cr
    n_iter = 1000
    rng = np.random.default_rng(5)
    scores = rng.uniform(1.0, 2.0, n_iter)
    scores[[137, 612, 880]] = 0.5                      # three equal minima: the lowest index must win everywhere
    scores[3] = np.nan
    out = []
    lo, hi = cr.shard_range(n_iter, rank, world)
    loc = scores[lo:hi]
    ok = np.where(np.isnan(loc), np.inf, loc)
    b = int(np.argmin(ok)); s = float(ok[b])
    out.append(cr.agree_on_winner(lo + b, s, n_iter, dist.group.WORLD, torch.device("cpu")))
    # a rank without any finite score
    if rank == 0:
        out.append(cr.agree_on_winner(-1, float("inf"), n_iter, dist.group.WORLD, torch.device("cpu")))
    else:
        out.append(cr.agree_on_winner(lo + b, s, n_iter, dist.group.WORLD, torch.device("cpu")))
    # nobody has one
    out.append(cr.agree_on_winner(-1, float("inf"), n_iter, dist.group.WORLD, torch.device("cpu")))
    return out


def test_ransac_winner_agreement_two_ranks():
    r0, r1 = _run(_winner_case)
    assert r0 == r1
    assert r0[0] == (137, 0.5)
    assert r0[1] == (612, 0.5)
    assert r0[2][0] == -1


def _ba_sum_case(rank, world):
    """Frame-sharded BA: every quantity the ranks exchange is a plain sum over frames (camera blocks, gradient,
    bone-length sums) — the packed-buffer all-reduce reproduces the single-process value.  Uses the oracle's
    Jacobian as the stand-in for the per-rank kernels."""
    from calib_b200 import scene
    from oracle import ba_oracle as bo
    sc = scene.make_scene(C=3, N=8, seed=77, score_mode="uniform", p_drop_2d=0.2)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    C, N, J = ss2d.shape
    per = N // world
    sl = slice(rank * per, (rank + 1) * per)
    X = sc.X_w[:, :J]

    def normal_blocks(frames, lam2_rows):
        a = (sc.K, sp2d[:, frames].reshape(C, -1, 2), ss2d[:, frames], sp3d[:, frames], ss3d[:, frames], bo.OP_BONE, C,
             ss2d[:, frames].shape[1], J, 1.0, 0.0 if not lam2_rows else 1.0)
        th = bo.to_theta(sc.R_init, sc.t_init, X[frames].reshape(-1, 3))
        f = bo.objfun(th, *a); Jm = bo.jacobian(th, *a).toarray()
        keep = slice(0, f.size - 12)                      # the 12 bone-variance rows couple all frames: excluded here
        Jc = Jm[keep, :6 * C]
        return np.concatenate([(Jc.T @ Jc).ravel(), Jc.T @ f[keep], [f[keep] @ f[keep]]])

    local = torch.from_numpy(normal_blocks(sl, False))
    dist.all_reduce(local)
    full = normal_blocks(slice(0, world * per), False)
    # bone-length means: global sum of lengths over frames, then the variance in a second pass (two-stage)
    L = np.linalg.norm(X[:, bo.OP_BONE[:, 0]] - X[:, bo.OP_BONE[:, 1]], axis=2)        # (N, 12)
    s1 = torch.from_numpy(np.concatenate([L[sl].sum(0), [per]])); dist.all_reduce(s1)
    mean = (s1[:12] / s1[12]).numpy()
    s2 = torch.from_numpy(((L[sl] - mean) ** 2).sum(0)); dist.all_reduce(s2)
    var = (s2 / s1[12]).numpy()
    return float(np.max(np.abs(local.numpy() - full)) / np.max(np.abs(full))), float(np.max(np.abs(var - L[:world * per].var(0))))


def test_ba_packed_sums_two_ranks():
    for e_blocks, e_var in _run(_ba_sum_case):
        assert e_blocks < 1e-12 and e_var < 1e-14


def _fake_solve(camid, K, R, t, sp2d, ss2d, sp3d, ss3d, lambda1, lambda2, **kw):
    """Stand-in for ba_main in the CPU tests of the sweep plumbing: a deterministic function of the problem."""
    return (R * lambda1, t + lambda2, np.full((2, 3), float(camid[0])))


def _batch_case(rank, world):
    from calib_b200 import ba
    n = 11
    probs = [(np.array([k]), None, np.full((2, 3, 3), float(k)), np.full((2, 3, 1), 0.5 * k), None, None, None, None)
             for k in range(n)]
    lazy = [(lambda p=p: p) for p in probs]                    # problems may be given as callables
    full = ba.ba_batch(lazy, 2.0, 3.0, n_streams=3, group=dist.group.WORLD, solve=_fake_solve)
    local = ba.ba_batch(probs, 2.0, 3.0, n_streams=1, group=dist.group.WORLD, gather=False, solve=_fake_solve)
    return ([float(r[2][0, 0]) for r in full], [float(r[0][0, 0, 0]) for r in full], [float(r[2][0, 0]) for r in local])


def test_ba_batch_sweep_two_ranks():
    """BASELINE configs[4] plumbing: problems dealt round-robin over the ranks, solved by worker threads, results
    gathered in index order on every rank (replicas only — no data-path collective)."""
    r0, r1 = _run(_batch_case)
    assert r0[0] == r1[0] == [float(k) for k in range(11)]
    assert r0[1] == [2.0 * k for k in range(11)]
    assert r0[2] == [float(k) for k in range(0, 11, 2)] and r1[2] == [float(k) for k in range(1, 11, 2)]


def test_ba_batch_plan_and_errors():
    from calib_b200 import ba
    for n in (0, 1, 5, 4096):
        for world in (1, 2, 8):
            plans = [ba.batch_plan(n, r, world) for r in range(world)]
            assert sorted(k for p in plans for k in p) == list(range(n))
            assert max(len(p) for p in plans) - min(len(p) for p in plans) <= 1

    def boom(*a, **k):
        raise ValueError("Initial guess is outside of provided bounds")
    with pytest.raises(ValueError, match="problem 0"):
        ba.ba_batch([(np.array([0]),) + (None,) * 7], 1.0, 1.0, solve=boom)
    with pytest.raises(calib_b200._lib.BaxError if not torch.cuda.is_available() else Exception):
        ba.ba_batch([(np.array([0]),) + (None,) * 7], 1.0, 1.0)     # no solve given: the CUDA path, no CPU fallback
