#!/bin/bash
# round-2 evidence run: GPU tests, smoke, default bench + reference arm, ncu launch list of the bench command, ncu --set full
# of k_rj_pipe / k_syrk96 / the LM kernels and of the RANSAC kernels (each ncu pass after the plain command exited 0)
mkdir -p gpurun_out
bash tools/gpu_validate.sh || exit 1
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 300 gpurun_out/bench_ref.json; echo
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_bench_default.csv python bench.py --steps 4 --warmup 3 --no-cpu --batch 4 > gpurun_out/ncu_bench.log 2>&1; echo "launch list rc=$?"
python tools/launch_summary.py gpurun_out/launches_bench_default.csv "ncu launch list of: python bench.py --steps 4 --warmup 3 --no-cpu --batch 4 (cold-cache, serialised per-launch times)" > gpurun_out/launches_bench_default_summary.txt; head -30 gpurun_out/launches_bench_default_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_rj_pipe" -s 2 -c 1 -o gpurun_out/k_rj_pipe_full python tools/prof_rj.py --workload cfg4shard --what rj --reps 3 > gpurun_out/ncu_rj.log 2>&1; tail -1 gpurun_out/ncu_rj.log
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_lin_fused|k_build_yhat|k_syrk96$|k_dense_solve1_coop|k_backsub|k_gp_bone" -c 8 -o gpurun_out/lm_kernels_full python tools/prof_rj.py --workload cfg4shard --what lm --reps 1 > gpurun_out/ncu_lm.log 2>&1; tail -1 gpurun_out/ncu_lm.log
python tools/ncu_summary.py gpurun_out/k_rj_pipe_full.ncu-rep gpurun_out/k_rj_pipe_ncu_full_summary.txt > /dev/null
python tools/ncu_summary.py gpurun_out/lm_kernels_full.ncu-rep gpurun_out/lm_kernels_ncu_full_summary.txt > /dev/null
ncu -i gpurun_out/k_rj_pipe_full.ncu-rep --page raw --csv > gpurun_out/k_rj_pipe_ncu_full_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_rsx_score|k_rsx_fit_warp|k_rsx_fit_deferred" -s 3 -c 3 -o gpurun_out/ransac_full python tools/bench_ransac.py > gpurun_out/ncu_ransac.log 2>&1; tail -1 gpurun_out/ncu_ransac.log
python tools/ncu_summary.py gpurun_out/ransac_full.ncu-rep gpurun_out/ransac_ncu_full_summary.txt > /dev/null
grep -E "kernel:|time_duration|fp64_cycles|dram_throughput|dram__bytes|warps_active|issue_active" gpurun_out/k_rj_pipe_ncu_full_summary.txt gpurun_out/lm_kernels_ncu_full_summary.txt gpurun_out/ransac_ncu_full_summary.txt | cut -c1-160
rm -f gpurun_out/lm_kernels_full.ncu-rep gpurun_out/ransac_full.ncu-rep gpurun_out/fit_lock*.ncu-rep
