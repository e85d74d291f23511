#!/bin/bash
# round-2 opening pass: round-end checks on the restored tree, FP64 peak record, compute-sanitizer on smoke
mkdir -p gpurun_out
bash tools/gpu_validate.sh
python tools/fp64_peak_json.py gpurun_out/fp64_peak.json
for tool in memcheck racecheck synccheck; do
  timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python __graft_entry__.py smoke > gpurun_out/sanitizer_${tool}_smoke.txt 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke ok" gpurun_out/sanitizer_${tool}_smoke.txt | tail -3
done
