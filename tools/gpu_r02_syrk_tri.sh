#!/bin/bash
# triangular diagonal tiles in k_syrk96: parity of the many-camera solves first, then the timing, then everything
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ba.py -q -x -m gpu -k "many_cameras or cooperative or full_size or sharded" > gpurun_out/syrk_tri_tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/syrk_tri_tests.log
for wl in cfg4shard cfg3; do timeout 300 python tools/time_syrk.py $wl 2>&1 | tail -n 1; done | tee gpurun_out/syrk_tri_time.log
bash tools/gpu_validate.sh
