#!/bin/bash
# Positional wrapper of calib_linear.py with the reference's 8-argument calling convention.
if [ $# != 8 ]; then
    echo
    echo "$0"
    echo "[Usage] PREFIX AID PID GID TARGET FRAME_SKIP DATASET OBS_MASK"
    echo "[e.g.] sh ./calib_linear.sh ./data/A023_P102_G003 23 102 3 noise_3_0 15 SynADL false"
    exit
fi
HERE="$(cd "$(dirname "$0")" && pwd)"
python3 "$HERE/calib_linear.py" --prefix "$1" --aid "$2" --pid "$3" --gid "$4" --target "$5" \
    --frame_skip "$6" --dataset "$7" --obs_mask "$8"
