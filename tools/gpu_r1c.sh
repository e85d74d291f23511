#!/bin/bash
# GPU pass: all GPU tests (BA + RANSAC), smoke, default bench, cfg1 bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
