#!/bin/bash
mkdir -p gpurun_out
python tools/prof_rj.py --what rj --reps 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_rj -s 1 -c 2 -o gpurun_out/prof_rj_bulk python tools/prof_rj.py --what rj --reps 3 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
ls -la gpurun_out
