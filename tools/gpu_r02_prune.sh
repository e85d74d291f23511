#!/bin/bash
# exact bail-out in the RANSAC score pass: tests, then the RANSAC leg with both modes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ransac.py -q -x -m gpu > gpurun_out/prune_tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/prune_tests.log
timeout 600 python tools/bench_ransac.py 2>&1 | tail -n 3 | cut -c1-1500 | tee gpurun_out/prune_bench.log
