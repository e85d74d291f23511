#!/bin/bash
mkdir -p gpurun_out
python tools/prof_ba_main.py 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/sweep_streams.py 64 > gpurun_out/sweep_streams.jsonl 2> gpurun_out/sweep_streams.err; echo "sweep rc=$?"; cat gpurun_out/sweep_streams.jsonl; tail -3 gpurun_out/sweep_streams.err
