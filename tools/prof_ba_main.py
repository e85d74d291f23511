#!/usr/bin/env python3
"""Where a cfg1-sized ba_main call spends its wall time (single stream): set-up vs iterations vs read-back."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba, scene  # noqa: E402

sc = scene.cfg1(seed=7000, N=1000)
sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
C, N, J = ss2d.shape
acc = {}


def tick(name, t0):
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0
    return time.perf_counter()


for rep in range(12):
    if rep == 2:
        acc.clear()
    t = time.perf_counter()
    prob = ba.BAProblem(sc.K, np.asarray(sp2d).reshape(C, N * J, 2), ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 1.0)
    t = tick("BAProblem()", t)
    x_all = prob.triangulate(sc.R_init, sc.t_init).reshape(N * J, 3)
    t = tick("triangulate", t)
    th0 = ba.to_theta(sc.R_init, sc.t_init, x_all)
    ok = np.all(np.isfinite(th0))
    t = tick("to_theta", t)
    res = ba.trf_schur(prob, th0)
    t = tick("trf_schur", t)
    th = res["x"].cpu().numpy()
    t = tick("x.cpu", t)
    prob.close()
    t = tick("close", t)
    ba.from_theta(th, C)
    t = tick("from_theta", t)
h = None
acc2 = {}
for rep in range(12):
    if rep == 2:
        acc2.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ups = [torch.as_tensor(a).to("cuda", torch.float64) for a in (sc.K, sp2d, ss2d, sp3d, ss3d)]
    torch.cuda.synchronize()
    acc2["uploads alone"] = acc2.get("uploads alone", 0.0) + time.perf_counter() - t0
    t0 = time.perf_counter()
    prob = ba.BAProblem(sc.K, np.asarray(sp2d).reshape(C, N * J, 2), ss2d, sp3d, ss3d, ba.OP_BONE, 1.0, 1.0, handle=h)
    torch.cuda.synchronize()
    acc2["BAProblem(handle=reused)"] = acc2.get("BAProblem(handle=reused)", 0.0) + time.perf_counter() - t0
    h = prob.detach()
print({k: round(v / 10 * 1e3, 3) for k, v in acc2.items()})
print({k: round(v / 10 * 1e3, 3) for k, v in acc.items()}, "ms per call; iterations", res["iterations"])
