"""Native batch sweep: fixed set-up cost per call vs per-problem rate (T = T0 + n / rate)."""
import time, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import calib_b200
from calib_b200 import ba, scene
scenes = []
for d in range(16):
    sc = scene.cfg1(seed=7000 + d, N=1000)
    scenes.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
ba.ba_batch([scenes[k % 16] for k in range(48)], 1.0, 1.0, n_streams=12)
for n in (12, 48, 192, 768, 1536):
    probs = [scenes[k % 16] for k in range(n)]
    t0 = time.perf_counter(); ba.ba_batch(probs, 1.0, 1.0, n_streams=12); dt = time.perf_counter() - t0
    print(n, "problems: %.3f s  (%.0f problems/s)" % (dt, n / dt), flush=True)
