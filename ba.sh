#!/bin/bash
# Positional wrapper of ba.py with the reference's 11-argument calling convention.
if [ $# != 11 ]; then
    echo
    echo "$0"
    echo "[Usage] PREFIX AID PID GID FRAME_SKIP LAMBDA1 LAMBDA2 TARGET DATASET OBS_MASK SAVE_OBS_MASK"
    echo "[e.g.] sh ./ba.sh ./data/A023_P102_G003 23 102 3 20 1. 1. linear_3_0 SynADL false false"
    exit
fi
HERE="$(cd "$(dirname "$0")" && pwd)"
python3 "$HERE/ba.py" --prefix "$1" --aid "$2" --pid "$3" --gid "$4" --frame_skip "$5" \
    --ba_lambda1 "$6" --ba_lambda2 "$7" --target "$8" --dataset "$9" \
    --obs_mask "${10}" --th_obs_mask 20 --save_obs_mask "${11}"
