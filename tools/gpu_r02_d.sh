#!/bin/bash
# native batch sweep: bit-identity test, throughput vs worker count, vs the round-1 Python-threaded path
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ba.py -m gpu -q -x -k "batch or converged or dropin" 2>&1 | tail -4
python - <<'PY'
import time, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
import calib_b200
from calib_b200 import ba, scene
scenes = []
for d in range(32):
    sc = scene.cfg1(seed=7000 + d, N=1000)
    scenes.append((sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc)))
probs = [scenes[k % 32] for k in range(512)]
ba.ba_batch(probs[:24], 1.0, 1.0, n_streams=12)
rows = []
for nw in (1, 4, 8, 12, 16, 24, 32):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = ba.ba_batch(probs, 1.0, 1.0, n_streams=nw, return_info=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    its = sum(r[3]["iterations"] for r in out)
    rows.append({"workers": nw, "problems_per_s": len(probs) / dt, "lm_iterations": its, "obs_per_s_per_iter": 68000 * its / dt})
    print(rows[-1], flush=True)
os.environ["BAX_BATCH_HOST"] = "1"
torch.cuda.synchronize(); t0 = time.perf_counter()
out = ba.ba_batch(probs[:128], 1.0, 1.0, n_streams=12, return_info=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
rows.append({"python_threads_round1_path": 12, "problems_per_s": 128 / dt})
print(rows[-1])
json.dump(rows, open("gpurun_out/batch_native_workers.json", "w"), indent=1)
PY
