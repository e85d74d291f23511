#!/bin/bash
# k_syrk96 experiment: variant 0 = quadrant kernel of round 2, 1 = triangular diagonal tiles with the extras branch inside the
# loop, 2 = templated + balanced warps; HOMOG = CTAs i and i+148 of one kind; NS = K cuts of (off-diagonal, diagonal) tiles
mkdir -p gpurun_out
run() { env "$@" timeout 300 python tools/time_syrk.py $WL 2>&1 | tail -n 1; }
{
WL=cfg4shard@16
for v in 0 1 2; do run BAX_SYRK_VARIANT=$v; done
WL=cfg4shard
run BAX_SYRK_VARIANT=0 BAX_SYRK_NS=29,29
for v in 1 2; do for hg in 0 1; do for ns in 32,26 33,24 32,25; do run BAX_SYRK_VARIANT=$v BAX_SYRK_HOMOG=$hg BAX_SYRK_NS=$ns; done; done; done
WL=cfg3
run BAX_SYRK_VARIANT=0 BAX_SYRK_NS=98,98
for v in 1 2; do for ns in 114,91 118,89 110,93; do run BAX_SYRK_VARIANT=$v BAX_SYRK_NS=$ns; done; done
} | tee gpurun_out/syrk_variants.log
