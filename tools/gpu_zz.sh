#!/bin/bash
# the no-alloc entry point's test alone in a fresh process, then the whole GPU suite (which collects it last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x -m gpu > gpurun_out/zz_alone.log 2>&1; echo "alone rc=$?"; tail -n 4 gpurun_out/zz_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest_gpu.log
