#!/bin/bash
for wd in 26 30 33 36 40; do BAX_SYRK_WD=$wd python tools/time_syrk.py cfg4shard 2>&1 | tail -1; done
BAX_SYRK_WD=26 python tools/time_syrk.py cfg3 2>&1 | tail -1
BAX_SYRK_WD=36 python tools/time_syrk.py cfg3 2>&1 | tail -1
