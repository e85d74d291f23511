#!/bin/bash
# Host build of the RANSAC device numerics for the CPU-only tests (test infrastructure, not product code).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
if [ ! -f librsx_hostemu.so ] || [ rsx_hostemu.cu -nt librsx_hostemu.so ] || [ ../../extrinsic-camera-calibration-from-a-moving-person_b200/csrc/rsx_fit.cuh -nt librsx_hostemu.so ] || [ ../../extrinsic-camera-calibration-from-a-moving-person_b200/csrc/rsx_internal.cuh -nt librsx_hostemu.so ] || [ ../../extrinsic-camera-calibration-from-a-moving-person_b200/csrc/rsx_arrow.cuh -nt librsx_hostemu.so ]; then
  $NVCC -O2 -std=c++17 -Wno-deprecated-gpu-targets -shared -Xcompiler -fPIC -o librsx_hostemu.so rsx_hostemu.cu
fi
echo "built $(pwd)/librsx_hostemu.so"
