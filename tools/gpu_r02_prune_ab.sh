#!/bin/bash
# rays per thread between two bail-out checks of the pruned score pass: 1, 2, 4 (default build), 8
P=extrinsic-camera-calibration-from-a-moving-person_b200
cp $P/libbax.so $P/libbax_clean.so
for v in pr1 pr2 clean pr8; do
  cp $P/libbax_$v.so $P/libbax.so
  timeout 600 python tools/bench_ransac.py 2>&1 | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', 'exact %.1f ms' % d['ms_per_step'], 'pruned %.2f ms' % d['pruned']['ms_per_step'], d['pruned']['dropped_on_this_rank'], d['pruned']['same_winner'])"
done | tee gpurun_out/prune_ab.log
cp $P/libbax_clean.so $P/libbax.so
