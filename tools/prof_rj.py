#!/usr/bin/env python
"""Short driver for ncu captures: binds one BA shard and launches each hot kernel a few times.

    python tools/prof_rj.py [--workload cfg4shard|cfg3|cfg1] [--reps 3] [--what rj|lm|all]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg4shard")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--what", default="all")
    a = ap.parse_args()
    import torch
    import calib_b200  # noqa: F401
    from calib_b200 import ba
    sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload(a.workload, 1000)
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = prob._dev_theta(ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3)))
    if a.what in ("rj", "all"):
        out = prob.residual_jacobian(th, with_index=False)
        for _ in range(a.reps):
            prob.residual_jacobian(th, with_index=False, out=out)
        del out
    if a.what in ("lm", "all"):
        sol = ba.TRFSolver(prob, th, ftol=0.0, xtol=0.0, gtol=0.0)
        for _ in range(a.reps):
            sol.iterate()
    torch.cuda.synchronize()
    print("prof ok", prob.n_vis)
    prob.close()


if __name__ == "__main__":
    main()
