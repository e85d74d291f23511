#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --workload cfg4shard --no-cpu --batch 0 > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_x.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_x.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g ms/step %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "rj", d["roofline"]["frac"], d["lm"])
PY
