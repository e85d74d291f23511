#!/usr/bin/env python3
"""Time the damped solve's kernel families at the cfg4 shard (BaxTimer split) — env knobs apply."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg4shard"
if "@" in wl:                                    # cfg4shard@16: the same workload with 16 cameras
    base, cams = wl.split("@")
    bench.WORKLOADS[wl] = dict(bench.WORKLOADS[base], C=int(cams))
sc, (sp2d, ss2d, sp3d, ss3d), w = bench.make_workload(wl, 1000)
prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"])
x0 = prob.triangulate(sc.R_init, sc.t_init)
th = prob._dev_theta(ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3)))
out = prob.linearize(th)
alpha = 0.01 * out[2] / out[3]
for _ in range(3):
    prob.trial(alpha, False)
prob.timing(True)
for _ in range(10):
    prob.trial(alpha, False)
torch.cuda.synchronize()
print(wl, {k: v for k, v in os.environ.items() if k.startswith("BAX_SYRK")}, {k: round(v[0] / 10, 4) for k, v in prob.timing_read().items() if v[1]}, flush=True)
prob.close()
