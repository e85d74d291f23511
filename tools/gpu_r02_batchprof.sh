#!/bin/bash
# sweep throughput: interpreter work pipelined with the native solve in chunks (BAX_BATCH_CHUNK; 0 = one call)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ba.py -q -x -m gpu -k "batch" > gpurun_out/batch_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/batch_tests.log
{ for c in 0 32 64 96 128 256; do BAX_BATCH_CHUNK=$c WORKERS=12 timeout 600 python tools/prof_batch.py 2>&1 | grep -E "FRAMES" | sed "s/^/chunk=$c /"; done
  WORKERS=12 timeout 600 python tools/prof_batch.py 2>&1 | grep -E "FRAMES" | sed "s/^/chunk=default /"; } | tee gpurun_out/batch_chunks.log
