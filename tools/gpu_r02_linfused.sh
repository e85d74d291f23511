#!/bin/bash
# fused camera + point linearisation (k_lin_fused) against k_lin_cam_mma + k_lin_pt
mkdir -p gpurun_out
{
for v in "4 1" "5 1" "5 2" "4 1" "5 1"; do set -- $v
  BAX_LINCAM=$1 BAX_LINFUSED_R=$2 timeout 600 python tools/time_lin.py cfg4shard cfg3 cfg1 2>&1 | grep linearize | sed "s/^/R=$2 /"
done
} | tee gpurun_out/linfused.log
timeout 900 python -m pytest tests/test_gpu_ba.py -q -x -m gpu > gpurun_out/linfused_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/linfused_tests.log
BAX_LINFUSED_R=2 timeout 900 python -m pytest tests/test_gpu_ba.py -q -x -m gpu > gpurun_out/linfused_tests_r2.log 2>&1; echo "tests R=2 rc=$?"; tail -n 3 gpurun_out/linfused_tests_r2.log
