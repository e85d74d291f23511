#!/bin/bash
# ncu --set full of one launch of each RANSAC kernel (fit, score) in the cfg2 leg
mkdir -p gpurun_out
python tools/bench_ransac.py > gpurun_out/ransac_plain.json 2> gpurun_out/ransac_plain.err || { tail gpurun_out/ransac_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k "regex:k_rsx_score|k_rsx_fit_warp" -s 4 -c 2 -o gpurun_out/ransac_full python tools/bench_ransac.py > gpurun_out/ncu_ransac.log 2>&1
tail -2 gpurun_out/ncu_ransac.log
python tools/ncu_summary.py gpurun_out/ransac_full.ncu-rep gpurun_out/ransac_ncu_full_summary.txt > /dev/null; cat gpurun_out/ransac_ncu_full_summary.txt
