#!/usr/bin/env python3
"""Frame-sharded BA over N GPUs (torchrun, NCCL) against the same scene solved on one GPU.

Checks, on a 6-camera scene with partial visibility whose frames are split in contiguous blocks over the ranks:
  * cost, gradient norms and one damped step (||p||^2, g.p, p^T (H + alpha I)^-1 p) of the sharded problem equal the
    single-GPU values (1e-11), the lazily computed phi' (bax_trial_phi_prime + BAX_BUF_RHS2_HEAD all-reduce) equals the
    eager one;
  * a complete trust-region solve starts on the same (alpha, cost) sequence and ends, after fixing the gauge (camera 0
    at the origin, unit baseline to camera 1), on the same cameras (1e-6) at the same cost (1e-9).
Prints one JSON line on rank 0; exit code 1 on mismatch."""
import json
import os
import sys

import ctypes

import numpy as np

C_int64 = ctypes.c_int64

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba, scene  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    per = 400
    N = per * world
    sc = scene.make_scene(C=6, N=N, seed=77, score_mode="uniform", p_drop_2d=0.25, p_drop_3d=0.1, sigma_px=2.0)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    C, _, J = ss2d.shape
    sl = slice(rank * per, (rank + 1) * per)
    lam1, lam2 = 2.0, 30.0
    full = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, lam1, lam2)
    x0 = full.triangulate(sc.R_init, sc.t_init)                       # (N, J, 3)
    th_full = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    th_loc = ba.to_theta(sc.R_init, sc.t_init, x0[sl].reshape(-1, 3))
    shard = ba.BAProblem(sc.K, sp2d[:, sl], ss2d[:, sl], sp3d[:, sl], ss3d[:, sl], ba.OP_BONE, lam1, lam2,
                         group=dist.group.WORLD, n_frames_global=N)
    out = {"world": world, "frames": N}
    worst = 0.0

    def rel(a, b):
        return abs(a - b) / max(abs(b), 1e-300)
    lf, ls = full.linearize(th_full), shard.linearize(th_loc)
    out["linearize_rel"] = [rel(a, b) for a, b in zip(ls, lf)]
    worst = max(worst, *out["linearize_rel"])
    alpha = 0.01 * lf[2] / lf[3]
    tf = full.trial(alpha, True)
    ts = shard.trial(alpha, True)
    out["trial_rel"] = [rel(a, b) for a, b in zip(ts, tf)]
    worst = max(worst, *out["trial_rel"])
    ts0 = shard.trial(alpha, False)
    lazy = shard.trial_phi_prime()
    out["lazy_phi_rel"] = rel(lazy, ts[2])
    out["lazy_pn2_rel"] = rel(ts0[0], ts[0])
    worst = max(worst, out["lazy_phi_rel"], out["lazy_pn2_rel"])
    # camera part of the step identical, point part = the rank's slice
    pf, ps = full.step.cpu().numpy(), shard.step.cpu().numpy()
    out["step_cam_rel"] = float(np.max(np.abs(pf[:6 * C] - ps[:6 * C])) / np.max(np.abs(pf[:6 * C])))
    pts_f = pf[6 * C:].reshape(N, J, 3)[sl].reshape(-1)
    out["step_pts_rel"] = float(np.max(np.abs(pts_f - ps[6 * C:])) / np.max(np.abs(pts_f)))
    worst = max(worst, out["step_cam_rel"], out["step_pts_rel"])
    # complete solves
    s_full = ba.TRFSolver(full, th_full, ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=30, native=False)
    s_shard = ba.TRFSolver(shard, th_loc, ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=30, native=False)
    assert not s_shard.native
    while not s_full.done():
        s_full.iterate()
    while not s_shard.done():
        s_shard.iterate()
    # the same sharded solve through the NATIVE driver: collectives issued inside libbax (bax_allreduce_attach), one C
    # call per outer iteration — must walk the very same path as the host-language loop above (same statements, same
    # collectives in the same order: NCCL sums are deterministic for a fixed communicator)
    out["native_comm"] = shard.native_comm is not None
    if shard.native_comm is not None:
        s_nat = ba.TRFSolver(shard, th_loc, ftol=1e-10, xtol=1e-12, gtol=1e-12, max_nfev=30)
        assert s_nat.native
        while not s_nat.done():
            s_nat.iterate()
        hn = s_nat.history
        out["native_nfev"] = s_nat.nfev
        out["native_vs_host_cost_rel"] = rel(s_nat.cost, s_shard.cost)
        out["native_vs_host_x_abs"] = float(torch.max(torch.abs(s_nat.x - s_shard.x)).item())
        out["native_vs_host_history_rel"] = max([rel(a["cost_new"], b["cost_new"]) for a, b in zip(hn, s_shard.history)] + [0.0])
        st = (C_int64 * 2)()
        shard.L.bax_comm_stats(shard.h, st)
        out["native_collectives"] = int(st[1])
        worst = max(worst, out["native_vs_host_cost_rel"], out["native_vs_host_history_rel"],
                    0.0 if s_nat.nfev == s_shard.nfev else 1.0)
    out["nfev"] = [s_full.nfev, s_shard.nfev]
    hf, hs = s_full.history, s_shard.history
    out["history_len"] = [len(hf), len(hs)]
    # A step taken at the alpha floor (1e-13 x the Hessian scale) is dominated, in the 7-dimensional gauge null space of the
    # exact Jacobian, by round-off / alpha — it differs between summation orders (shard counts) while staying a valid
    # trust-region step; the cost along the path then agrees to ~1e-5 only (second-order effect of the gauge motion).
    # What must agree is the end point after fixing the gauge, and the cost there.
    k = min(4, len(hf), len(hs))
    hist = max([rel(a["cost_new"], b["cost_new"]) for a, b in zip(hs[:k], hf[:k])] +
               [rel(a["alpha"], b["alpha"]) for a, b in zip(hs[:k], hf[:k])])
    out["history_rel_first4"] = hist
    out["hist_full"] = [(a["alpha"], a["Delta"], a["cost_new"], a["n_it"]) for a in hf[:6]]
    out["hist_shard"] = [(a["alpha"], a["Delta"], a["cost_new"], a["n_it"]) for a in hs[:6]]
    xf, xs = s_full.x.cpu().numpy(), s_shard.x.cpu().numpy()

    def gauge(theta):
        R, tt, _ = ba.from_theta(theta, C)
        tt = tt.reshape(C, 3)
        Rn = np.array([R[i] @ R[0].T for i in range(C)])
        tn = np.array([tt[i] - Rn[i] @ tt[0] for i in range(C)])
        return Rn, tn / np.linalg.norm(tn[1])
    (Rf, tf_), (Rs, ts_) = gauge(xf), gauge(xs)
    out["R_abs_gauge_fixed"] = float(np.max(np.abs(Rf - Rs)))
    out["t_abs_gauge_fixed"] = float(np.max(np.abs(tf_ - ts_)))
    out["cost_rel"] = rel(s_shard.cost, s_full.cost)
    ok = (worst < 1e-9 and hist < 1e-3 and out["R_abs_gauge_fixed"] < 1e-6 and out["t_abs_gauge_fixed"] < 1e-6
          and out["cost_rel"] < 1e-9)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    out["ok"] = bool(flag.item() == 0)
    if rank == 0:
        print(json.dumps(out))
    full.close(); shard.close()
    dist.destroy_process_group()
    sys.exit(0 if out["ok"] else 1)


if __name__ == "__main__":
    main()
