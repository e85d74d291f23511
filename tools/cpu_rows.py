#!/usr/bin/env python3
"""CPU rows that stand beside the GPU numbers, timed on this box's host cores:

  literal : SciPy least_squares(trf, dense 2-point finite differences, dense SVD) on the oracle port of ba.objfun —
            exactly what reference ba.py:187-206 runs — at frame counts small enough to execute, giving measured
            seconds per trust-region iteration and observations/s per iteration;
  strong  : NOT the reference: the same objective with the oracle's analytic sparse Jacobian through
            least_squares(jac=..., tr_solver='lsmr') at the cfg1 size — the honest "good CPU implementation" row.

Test/bench infrastructure (imports oracle/); prints one JSON object."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from scipy.optimize import least_squares
    import calib_b200  # noqa: F401
    from calib_b200 import scene
    from oracle import ba_oracle as bo
    out = {"host_cpus": os.cpu_count(), "literal": [], "strong": None}
    budget = float(os.environ.get("CPU_ROWS_BUDGET_S", "60"))
    t_start = time.perf_counter()
    for N in (5, 10, 20, 40):
        if time.perf_counter() - t_start > budget:
            break
        sc = scene.cfg1(seed=300 + N, N=N)
        sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
        C, _, J = ss2d.shape
        x0 = bo.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init.reshape(C, 3, 1), ss2d > 0)
        th0 = bo.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 1.0, 1.0)
        t0 = time.perf_counter()
        res = least_squares(bo.objfun, th0, method="trf", ftol=1e-4, args=args)
        dt = time.perf_counter() - t0
        n_obs = int((ss2d > 0).sum())
        out["literal"].append({"frames": N, "n": int(th0.size), "m": int(res.fun.size), "observations": n_obs,
                               "seconds": dt, "iterations": int(res.njev), "nfev": int(res.nfev),
                               "s_per_iteration": dt / res.njev, "obs_per_s_per_iteration": n_obs * res.njev / dt,
                               "status": int(res.status)})
    # strong CPU row at cfg1
    sc = scene.cfg1(seed=0, N=int(os.environ.get("CPU_ROWS_STRONG_N", "1000")))
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    C, N, J = ss2d.shape
    t0 = time.perf_counter()
    x0 = bo.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init.reshape(C, 3, 1), ss2d > 0)
    t_tri = time.perf_counter() - t0
    th0 = bo.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 1.0, 1.0)
    t0 = time.perf_counter()
    res = least_squares(bo.objfun, th0, jac=bo.jacobian, method="trf", tr_solver="lsmr", ftol=1e-4, args=args,
                        max_nfev=int(os.environ.get("CPU_ROWS_STRONG_NFEV", "12")))
    dt = time.perf_counter() - t0
    n_obs = int((ss2d > 0).sum())
    out["strong"] = {"frames": N, "n": int(th0.size), "observations": n_obs, "seconds": dt, "iterations": int(res.njev),
                     "nfev": int(res.nfev), "s_per_iteration": dt / res.njev,
                     "obs_per_s_per_iteration": n_obs * res.njev / dt, "status": int(res.status),
                     "triangulation_s": t_tri,
                     "note": "analytic sparse Jacobian + LSMR trust-region solver; not what the reference runs"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
