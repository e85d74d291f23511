import time, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import calib_b200
from calib_b200 import ba, scene
scenes = []
for d in range(16):
    sc = scene.cfg1(seed=7000 + d, N=1000)
    scenes.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
probs = [scenes[k % 16] for k in range(384)]
ba.ba_batch(probs[:24], 1.0, 1.0, n_streams=12)
os.environ["BAX_BATCH_PROFILE"] = "1"
for nw in (1, 12):
    t0 = time.perf_counter(); out = ba.ba_batch(probs, 1.0, 1.0, n_streams=nw, return_info=True); dt = time.perf_counter() - t0
    print(nw, "workers:", len(probs) / dt, "problems/s; iterations per problem", sum(r[3]["iterations"] for r in out) / len(out), "nfev", sum(r[3]["nfev"] for r in out) / len(out))
