#!/bin/bash
mkdir -p gpurun_out
python tools/prof_rj.py --workload cfg4shard --what lm --reps 1 > gpurun_out/plain_lm.log 2>&1 || { tail gpurun_out/plain_lm.log; exit 1; }
ncu --set full --clock-control none --import-source on -k "regex:k_lin_cam_mma|k_build_yhat|k_backsub|k_lin_pt|k_gemv_last" -c 5 -o gpurun_out/lm_kernels_full python tools/prof_rj.py --workload cfg4shard --what lm --reps 1 > gpurun_out/ncu_lm_full.log 2>&1
tail -2 gpurun_out/ncu_lm_full.log
python tools/ncu_summary.py gpurun_out/lm_kernels_full.ncu-rep gpurun_out/lm_kernels_ncu_full_summary.txt > /dev/null; cat gpurun_out/lm_kernels_ncu_full_summary.txt
