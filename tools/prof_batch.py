"""Native batch sweep: throughput vs worker count (and CUDA_DEVICE_MAX_CONNECTIONS from the environment)."""
import time, os, sys, json
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import calib_b200
from calib_b200 import ba, scene
scenes = []
for d in range(16):
    sc = scene.cfg1(seed=7000 + d, N=int(os.environ.get('FRAMES', '1000')))
    scenes.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
probs = [scenes[k % 16] for k in range(512)]
ba.ba_batch(probs[:48], 1.0, 1.0, n_streams=24)
rows = []
for nw in [int(x) for x in os.environ.get("WORKERS", "1,12").split(",")]:
    t0 = time.perf_counter(); out = ba.ba_batch(probs, 1.0, 1.0, n_streams=nw, return_info=True); dt = time.perf_counter() - t0
    rows.append((nw, round(len(probs) / dt, 1)))
print("FRAMES", os.environ.get("FRAMES"), "CUDA_DEVICE_MAX_CONNECTIONS", os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), rows)
