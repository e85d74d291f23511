#!/bin/bash
# Positional wrapper of calib_ransac.py with the reference's 7-argument calling convention.
# RANSAC_NITER / RANSAC_SEED may be exported to override the parser defaults (100 / 0).
if [ $# != 7 ]; then
    echo
    echo "$0"
    echo "[Usage] PREFIX AID PID GID TARGET FRAME_SKIP DATASET"
    echo "[e.g.] sh ./calib_ransac.sh ./data/A023_P102_G003 23 102 3 noise_3_0 15 SynADL"
    exit
fi
HERE="$(cd "$(dirname "$0")" && pwd)"
python3 "$HERE/calib_ransac.py" --prefix "$1" --aid "$2" --pid "$3" --gid "$4" --target "$5" \
    --frame_skip "$6" --dataset "$7" ${RANSAC_NITER:+--ransac_niter $RANSAC_NITER} ${RANSAC_SEED:+--ransac_seed $RANSAC_SEED}
