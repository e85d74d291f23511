#!/usr/bin/env python3
"""Sweep of worker-stream counts for ba_batch on cfg1-shaped problems (BASELINE configs[4] sample)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import calib_b200  # noqa: E402,F401
from calib_b200 import ba, scene  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
probs = []
for k in range(n):
    sc = scene.cfg1(seed=7000 + k, N=1000)
    probs.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
import threading  # noqa: E402
durs = []
lock = threading.Lock()


def timed(*a, **k):
    t0 = time.perf_counter()
    r = ba.ba_main(*a, **k)
    with lock:
        durs.append(time.perf_counter() - t0)
    return r


ba.ba_batch(probs[:32], 1.0, 1.0, n_streams=32)
for rep in range(2):
    for ns in (1, 2, 4, 8, 12, 16, 24, 32):
        durs.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = ba.ba_batch(probs, 1.0, 1.0, n_streams=ns, return_info=True, solve=timed, verbose=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        it = sum(r[3]["iterations"] for r in out)
        d = sorted(durs)
        print(json.dumps({"streams": ns, "problems": n, "problems_per_s": n / dt, "lm_iterations": it,
                          "obs_per_s_per_lm_iteration": 68000.0 * it / dt, "call_ms_median": 1e3 * d[len(d) // 2],
                          "call_ms_max3": [round(1e3 * x, 2) for x in d[-3:]]}), flush=True)
