#!/bin/bash
# reductions / tiny kernels folded into their producers: tests, cfg1 iteration, sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu.log
{ timeout 300 python tools/time_lin.py cfg1 cfg4shard 2>&1 | grep linearize | cut -c1-90
  for i in 1 2; do WORKERS=1,12 timeout 600 python tools/prof_batch.py 2>&1 | grep FRAMES; done; } | tee gpurun_out/fold2.log
