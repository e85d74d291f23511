#!/bin/bash
# first GPU pass: tests, bench, ncu launch list, full capture of k_rj
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_cfg4shard.json 2> gpurun_out/bench_cfg4shard.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_cfg4shard.json
python bench.py --workload cfg1 --no-cpu > gpurun_out/bench_cfg1.json 2> gpurun_out/bench_cfg1.err
python bench.py --workload cfg3 --no-cpu > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python tools/prof_rj.py --what all --reps 2 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_lm.csv python tools/prof_rj.py --what all --reps 2 > gpurun_out/ncu1.log 2>&1
python tools/prof_rj.py --what rj --reps 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_rj -s 1 -c 2 -o gpurun_out/prof_rj python tools/prof_rj.py --what rj --reps 3 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
