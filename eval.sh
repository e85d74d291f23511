#!/bin/bash
# Positional wrapper of eval.py with the reference's 7-argument calling convention.
if [ $# != 7 ]; then
    echo
    echo "$0"
    echo "[Usage] PREFIX AID PID GID FRAME_SKIP TARGET DATASET"
    echo "[e.g.] sh ./eval.sh ./data/A023_P102_G003 23 102 3 15 linear_3_0 SynADL"
    exit
fi
HERE="$(cd "$(dirname "$0")" && pwd)"
echo "$7"
python3 "$HERE/eval.py" --prefix "$1" --aid "$2" --pid "$3" --gid "$4" --frame_skip "$5" --target "$6" --dataset "$7"
