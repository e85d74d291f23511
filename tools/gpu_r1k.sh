#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_ransac.py > gpurun_out/ransac_k.json 2> gpurun_out/ransac_k.err; echo "ransac rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/ransac_k.json")); print(d["value"], d["e2e"]["value"], d["kernel_ms_one_pass"], d["winner"])
PY
