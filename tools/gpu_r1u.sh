#!/bin/bash
mkdir -p gpurun_out
python tools/time_lin.py cfg4shard cfg3 cfg1 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
