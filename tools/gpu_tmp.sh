#!/bin/bash
mkdir -p gpurun_out
BAX_CHOL_WARP=0 python tools/time_syrk.py cfg4shard 2>&1 | tail -1
BAX_CHOL_WARP=1 python tools/time_syrk.py cfg4shard 2>&1 | tail -1
BAX_CHOL_WARP=1 BAX_DENSE_COOP=2 python tools/time_syrk.py cfg4shard 2>&1 | grep "dense coop" | tail -1
BAX_CHOL_WARP=0 python tools/time_syrk.py cfg3 2>&1 | tail -1
BAX_CHOL_WARP=1 python tools/time_syrk.py cfg3 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
