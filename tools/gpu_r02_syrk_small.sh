#!/bin/bash
# warp-per-unit DMMA SYRK for up to 48 rows (4-camera problems) against the 64 x 64 DFMA tile kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ba.py -q -x -m gpu > gpurun_out/syrk_small_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/syrk_small_tests.log
{
for v in 0 1 0 1; do
  BAX_SYRK_SMALL=$v timeout 300 python tools/time_syrk.py cfg1 2>&1 | tail -n 1
  BAX_SYRK_SMALL=$v timeout 300 python tools/time_lin.py cfg1 2>&1 | grep linearize | cut -c1-80
done
for v in 0 1; do BAX_SYRK_SMALL=$v timeout 600 python tools/prof_batch.py 2>&1 | tail -n 4; done
} | tee gpurun_out/syrk_small.log
