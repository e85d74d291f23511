#!/bin/bash
# 2-GPU bench (as the driver launches it) with the comm-timing pass
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print("N=$N value %.4g e2e %.4g ms/step %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), "rj", d["roofline"]["frac"], "ransac", d["ransac"]["value"], d["ransac"]["winner"])
print(json.dumps(d["comm"])); print(json.dumps(d["batch"])); print(json.dumps(d["lm"]))
PY
