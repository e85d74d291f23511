#!/bin/bash
mkdir -p gpurun_out
for v in 0 1 2 3 4; do BAX_LINCAM=$v python tools/time_lin.py cfg4shard cfg3 cfg1 2>&1 | tail -3; done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
