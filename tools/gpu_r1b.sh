#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_ransac.py -m gpu -x -q > gpurun_out/pytest_ransac.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_ransac.log
tail -30 gpurun_out/pytest_ransac.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --workload cfg1 --steps 5 > gpurun_out/bench_cfg1_b.json 2> gpurun_out/bench_cfg1_b.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_cfg1_b.json; tail -5 gpurun_out/bench_cfg1_b.err
