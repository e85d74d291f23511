#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ba.py -m gpu -q -x -k "full_size or batch" > gpurun_out/pytest_fullsize.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_fullsize.log
timeout 300 python tools/sweep_streams.py 64 > gpurun_out/sweep_streams.jsonl 2> gpurun_out/sweep_streams.err; echo "sweep rc=$?"; cat gpurun_out/sweep_streams.jsonl; tail -3 gpurun_out/sweep_streams.err
CPU_ROWS_BUDGET_S=45 timeout 400 python bench.py --impl cpu_rows > gpurun_out/cpu_rows.json 2> gpurun_out/cpu_rows.err; echo "cpu rows rc=$?"; cat gpurun_out/cpu_rows.json
