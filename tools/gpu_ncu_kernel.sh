#!/bin/bash
# ncu --set full of one kernel (regex) in the LM loop; args: regex skip count outname [workload]
mkdir -p gpurun_out
W=${5:-cfg4shard}
ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $3 -o gpurun_out/$4 python tools/prof_rj.py --workload $W --what lm --reps 1 > gpurun_out/ncu_$4.log 2>&1
tail -2 gpurun_out/ncu_$4.log
