#!/bin/bash
# ncu --set full of one launch of the RANSAC score kernel + SASS-level source page (joined to source lines by tools/ncu_source_lines.py)
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_rsx_score" -s 1 -c 1 -o gpurun_out/score_full python tools/bench_ransac.py > gpurun_out/ncu_score.log 2>&1
ncu -i gpurun_out/score_full.ncu-rep --page source --csv > gpurun_out/score_source.csv 2>/dev/null; wc -l gpurun_out/score_source.csv
python tools/ncu_summary.py gpurun_out/score_full.ncu-rep gpurun_out/score_summary.txt | head -30
