#!/bin/bash
# A/B in one run: chunk width of the SYRK pipeline (16 columns x 3 stages against 32 x 2)
mkdir -p gpurun_out
P=extrinsic-camera-calibration-from-a-moving-person_b200
run() { env "$@" timeout 300 python tools/time_syrk.py $WL 2>&1 | tail -n 1; }
cp $P/libbax.so $P/libbax_clean.so
{
for v in clean k32 clean k32; do
  cp $P/libbax_$v.so $P/libbax.so
  echo "== $v"; WL=cfg4shard; run A=1; WL=cfg3; run A=1; WL=cfg4shard@16; run A=1
done
cp $P/libbax_k32.so $P/libbax.so
timeout 600 python -m pytest tests/test_gpu_ba.py -q -x -m gpu -k "many_cameras or cooperative or full_size" 2>&1 | tail -n 2
} | tee gpurun_out/syrk_ab_k32.log
cp $P/libbax_clean.so $P/libbax.so
