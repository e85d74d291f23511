#!/bin/bash
# k_rj variants (incl. the persistent TMA pipeline 4/5) + HBM traffic-mix ceiling + parity of the new variant
mkdir -p gpurun_out
timeout 120 tools/micro/hbm_mix > gpurun_out/hbm_mix.jsonl 2>&1; echo "hbm_mix rc=$?"; cat gpurun_out/hbm_mix.jsonl
timeout 600 python tools/rj_variants.py --workload cfg4shard --variants 2,4,5 > gpurun_out/rj_variants_h.jsonl 2> gpurun_out/rj_variants_h.err; echo "variants rc=$?"; cat gpurun_out/rj_variants_h.jsonl; tail -3 gpurun_out/rj_variants_h.err
timeout 600 python tools/rj_variants.py --workload cfg3 --variants 2,4,5 >> gpurun_out/rj_variants_h.jsonl 2>> gpurun_out/rj_variants_h.err; echo "variants cfg3 rc=$?"; tail -3 gpurun_out/rj_variants_h.jsonl
timeout 600 python tools/rj_variants.py --workload cfg1 --variants 2,4,5 >> gpurun_out/rj_variants_h.jsonl 2>> gpurun_out/rj_variants_h.err; echo "variants cfg1 rc=$?"; tail -3 gpurun_out/rj_variants_h.jsonl
BAX_RJ_VARIANT=4 timeout 900 python -m pytest tests/test_gpu_ba.py -m gpu -q -x > gpurun_out/pytest_v4.log 2>&1; echo "pytest v4 rc=$?"; tail -3 gpurun_out/pytest_v4.log
