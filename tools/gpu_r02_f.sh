#!/bin/bash
# launch list of one cfg1-sized ba_main (what a problem of the cfg5 sweep costs on the device)
mkdir -p gpurun_out
cat > /tmp/one_cfg1.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import calib_b200
from calib_b200 import ba, scene
sc = scene.cfg1(seed=7000, N=1000)
p = (sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc))
ba.ba_batch([p, p], 1.0, 1.0, n_streams=1)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_cfg1_problem.csv python /tmp/one_cfg1.py > gpurun_out/ncu_cfg1.log 2>&1; echo "rc=$?"
python tools/launch_summary.py gpurun_out/launches_cfg1_problem.csv | head -60
