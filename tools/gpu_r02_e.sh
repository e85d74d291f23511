#!/bin/bash
# 2 GPUs: sharded parity (host-language loop vs native driver with in-library NCCL), then the 2-GPU bench
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/check_sharded.py > gpurun_out/sharded_parity_n2.json 2> gpurun_out/sharded_parity_n2.err; echo "check_sharded rc=$?"; tail -3 gpurun_out/sharded_parity_n2.err; cat gpurun_out/sharded_parity_n2.json | cut -c1-1500
bash tools/gpu_bench_multi.sh 2
BAX_NATIVE_COMM=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 30 --warmup 3 --batch 0 > gpurun_out/bench_n2_hostcomm.json 2> gpurun_out/bench_n2_hostcomm.err; echo "host-comm bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n2_hostcomm.json").read().strip().splitlines()[-1])
print("host-language collectives: value %.4g ms/step %.3f" % (d["value"], d["ms_per_step"]))
PY
