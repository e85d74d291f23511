#!/bin/bash
# ncu --set full of one launch of the RANSAC warp fit kernel (free-running and lock-step), summary + hottest source lines
mkdir -p gpurun_out
for mode in 0 1; do
  RSX_FIT_LOCKSTEP=$mode timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_rsx_fit_warp" -s 6 -c 1 -o gpurun_out/fit_lock$mode python tools/bench_ransac.py > gpurun_out/ncu_fit$mode.log 2>&1
  python tools/ncu_summary.py gpurun_out/fit_lock$mode.ncu-rep gpurun_out/fit_lock${mode}_summary.txt > /dev/null; cat gpurun_out/fit_lock${mode}_summary.txt
done
ncu -i gpurun_out/fit_lock0.ncu-rep --page source --csv > gpurun_out/fit_lock0_source.csv 2>/dev/null; wc -l gpurun_out/fit_lock0_source.csv
