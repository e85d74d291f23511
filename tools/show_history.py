#!/usr/bin/env python3
"""Print the trust-region history (alpha, Delta, step norm, More' iterations, damped solves) of one complete solve."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba  # noqa: E402

for wl in sys.argv[1:] or ["cfg4shard"]:
    sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload(wl, 1000)
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    sol = ba.TRFSolver(prob, th)
    print(wl, "x_norm", sol.x_norm, "cost0", sol.cost, "h_scale", prob.h_scale, "floor", 1e-13 * prob.h_scale)
    while not sol.done():
        n0 = prob.n_trials
        sol.iterate()
        for h in sol.history[-1:]:
            print("  it", sol.iteration, {k: (f"{v:.4g}" if isinstance(v, float) else v) for k, v in h.items()},
                  "trials", prob.n_trials - n0, "status", sol.status)
    prob.close()
