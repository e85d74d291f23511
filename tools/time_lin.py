#!/usr/bin/env python3
"""Time one linearisation (and one damped solve) per workload; env BAX_LINCAM selects the k_lin_cam variant."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba  # noqa: E402

for wl in sys.argv[1:] or ["cfg4shard"]:
    sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload(wl, 1000)
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    th = prob._dev_theta(ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3)))
    out = prob.linearize(th)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        prob.linearize(th)
    e0.record()
    for _ in range(20):
        prob.linearize(th)
    e1.record(); torch.cuda.synchronize()
    t_lin = e0.elapsed_time(e1) / 20
    alpha = 0.01 * out[2] / out[3]
    prob.trial(alpha, False)
    e0.record()
    for _ in range(10):
        prob.trial(alpha, False)
    e1.record(); torch.cuda.synchronize()
    print(f"{wl} BAX_LINCAM={os.environ.get('BAX_LINCAM')} linearize {t_lin:.4f} ms  trial {e0.elapsed_time(e1) / 10:.4f} ms  cost {out[0]:.15e} g2 {out[2]:.15e}", flush=True)
    prob.close()
