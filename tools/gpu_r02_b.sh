#!/bin/bash
# GPU tests (new parity pins), RANSAC score-mode A/B, default bench with the new roofline_step / sharded_parity fields
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -k "not n40" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
for m in 0 1 2; do
  RSX_SCORE_MODE=$m timeout 600 python tools/bench_ransac.py > gpurun_out/ransac_mode$m.json 2> gpurun_out/ransac_mode$m.err; echo "mode $m rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/ransac_mode$m.json").read().strip().splitlines()[-1])
print("mode $m: %.4g hyp/s  ms/step %.2f" % (d["value"], d["ms_per_step"]), d["kernel_ms_one_pass"], "winner", d["winner"], d["inliers_2d"], d["inliers_3d"])
PY
done
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "rj", d["roofline"]["frac"], "ransac", d["ransac"]["value"])
print(json.dumps(d["roofline_step"])); print(json.dumps(d["sharded_parity"])); print(json.dumps(d["cpu_baseline"])[:1500])
PY
