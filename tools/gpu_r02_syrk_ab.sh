#!/bin/bash
# A/B in one run: unroll factors of the K loop (diagonal path, off-diagonal path): u44 = clean build
mkdir -p gpurun_out
P=extrinsic-camera-calibration-from-a-moving-person_b200
run() { env "$@" timeout 300 python tools/time_syrk.py $WL 2>&1 | tail -n 1; }
cp $P/libbax.so $P/libbax_clean.so
{
for v in clean u22 u24 u14 u11 exp clean; do
  cp $P/libbax_$v.so $P/libbax.so
  echo "== $v"; WL=cfg4shard; run A=1; run A=2; WL=cfg3; run A=1; WL=cfg4shard@16; run A=1
done
} | tee gpurun_out/syrk_ab.log
cp $P/libbax_clean.so $P/libbax.so
