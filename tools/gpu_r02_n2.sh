#!/bin/bash
# 2 GPUs with the final kernels: sharded solve against the single-GPU one (native collectives), then the bench as the driver launches it
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/check_sharded.py > gpurun_out/check_sharded_n2.json 2> gpurun_out/check_sharded_n2.err; echo "check_sharded rc=$?"; tail -c 600 gpurun_out/check_sharded_n2.json; echo
bash tools/gpu_bench_multi.sh 2
