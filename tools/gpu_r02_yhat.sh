#!/bin/bash
# k_build_yhat: cameras of a point dealt over 1, 2, 4, 8 CTAs ('build' bucket of the damped solve)
mkdir -p gpurun_out
{
for wl in cfg4shard cfg3 cfg4shard@16 cfg1; do for g in 0 1 2 4 8; do
  BAX_YHAT_GROUPS=$g timeout 300 python tools/time_syrk.py $wl 2>&1 | tail -n 1 | sed "s/^/groups=$g /"
done; done
} | tee gpurun_out/yhat_groups.log
timeout 900 python -m pytest tests/test_gpu_ba.py -q -x -m gpu > gpurun_out/yhat_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/yhat_tests.log
