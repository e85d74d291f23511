#!/bin/bash
# round-1 profiles: launch list of the default bench command, full captures of the top kernels
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python tools/prof_rj.py --what lm --reps 1 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_syrk96$" -s 1 -c 1 -o gpurun_out/prof_syrk_dmma python tools/prof_rj.py --what lm --reps 1 > gpurun_out/ncu_syrk.log 2>&1
echo "syrk rc=$?"
python tools/bench_ransac.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_rsx_score|k_rsx_fit_warp" -c 2 -o gpurun_out/prof_ransac python tools/bench_ransac.py > gpurun_out/ncu_ransac.log 2>&1
echo "ransac rc=$?"
ls -la gpurun_out | tail -8
