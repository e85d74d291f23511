#!/bin/bash
# RANSAC: GPU tests, timing of the fit layouts (RSX_FIT_MODE 1 = compact default, 5 = round-1 layout), ncu of both kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_ransac.py tests/test_dataio_cli.py -m gpu -q -x > gpurun_out/pytest_gpu_ransac.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_ransac.log
for m in 1 5; do
  RSX_FIT_MODE=$m timeout 600 python tools/bench_ransac.py > gpurun_out/ransac_fit$m.json 2> gpurun_out/ransac_fit$m.err; echo "fit mode $m rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/ransac_fit$m.json").read().strip().splitlines()[-1])
print("fit mode $m: %.4g hyp/s  ms/step %.2f" % (d["value"], d["ms_per_step"]), d["kernel_ms_one_pass"], "winner", d["winner"], d["inliers_2d"], d["inliers_3d"])
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_rsx_score|k_rsx_fit_warp|k_rsx_fit_deferred" -s 6 -c 3 -o gpurun_out/ransac_full python tools/bench_ransac.py > gpurun_out/ncu_ransac.log 2>&1
tail -2 gpurun_out/ncu_ransac.log
python tools/ncu_summary.py gpurun_out/ransac_full.ncu-rep gpurun_out/ransac_ncu_full_summary.txt > /dev/null; cat gpurun_out/ransac_ncu_full_summary.txt
