#!/bin/bash
mkdir -p gpurun_out
for w in cfg4shard cfg3 cfg1; do timeout 300 python tools/rj_variants.py --workload $w >> gpurun_out/rj_variants.log 2>&1; done
cat gpurun_out/rj_variants.log
