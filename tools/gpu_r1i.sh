#!/bin/bash
mkdir -p gpurun_out
for w in cfg4shard cfg3; do
timeout 600 python tools/rj_variants.py --workload $w --variants 2,4,5,6,7,8 --reps 30 >> gpurun_out/rj_variants_i.jsonl 2>> gpurun_out/rj_variants_i.err; echo "variants $w rc=$?"
done
cat gpurun_out/rj_variants_i.jsonl
