#!/bin/bash
# full GPU suite with the new default k_rj, default bench, ncu full capture of k_rj_pipe + launch list of the bench command
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_default.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_default.json"))
print("value %.4g e2e %.4g ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "rj", d["roofline"]["frac"], d["clocks"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], "ransac", d["ransac"]["value"])
PY
timeout 600 python tools/prof_rj.py --workload cfg4shard --what rj --reps 3 > gpurun_out/plain_rj.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rj_pipe -s 2 -c 1 -o gpurun_out/k_rj_pipe_full python tools/prof_rj.py --workload cfg4shard --what rj --reps 3 > gpurun_out/ncu_rj_pipe.log 2>&1
tail -2 gpurun_out/ncu_rj_pipe.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_bench_default.csv python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "launch list rc=$?"
ls -la gpurun_out | tail -12
