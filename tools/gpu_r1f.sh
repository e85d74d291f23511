#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ba.py -m gpu -q -x > gpurun_out/pytest_gpu_ba.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_ba.log
for w in cfg4shard cfg3 cfg1; do
timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$w.json"))
print("$w", "value %.4g e2e %.4g ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), d["lm"], "rj frac %.3f" % d["roofline"]["frac"])
PY
done
