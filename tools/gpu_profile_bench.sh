#!/bin/bash
# default bench, reference arm, launch list of the bench command, ncu full of the LM kernels (after prefetch)
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "rj", d["roofline"]["frac"], d["clocks"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "ransac", d["ransac"]["value"], "launches", d["gpu_launches"])
print(d["lm"]); print(d["batch"]["value"], d["batch"]["single_worker_problems_per_s_one_gpu"])
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 400 gpurun_out/bench_ref.json
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_bench_default.csv python bench.py --steps 4 --warmup 3 --no-cpu --batch 4 > gpurun_out/ncu_bench.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k "regex:k_lin_cam_mma|k_build_yhat|k_lin_pt|k_syrk96$|k_dense_solve1_coop" -c 6 -o gpurun_out/lm_kernels_full2 python tools/prof_rj.py --workload cfg4shard --what lm --reps 1 > gpurun_out/ncu_lm_full2.log 2>&1
python tools/ncu_summary.py gpurun_out/lm_kernels_full2.ncu-rep gpurun_out/lm_kernels_ncu_full_summary.txt > /dev/null; grep -E "kernel:|time_duration|fp64_cycles|dram_throughput|warps_active|stall" gpurun_out/lm_kernels_ncu_full_summary.txt | head -80
rm -f gpurun_out/lm_kernels_full.ncu-rep
