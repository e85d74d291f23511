#!/bin/bash
# ncu launch list (per-kernel durations) of two LM iterations; args: workload
W=${1:-cfg4shard}
mkdir -p gpurun_out
python tools/prof_rj.py --workload $W --what lm --reps 2 > gpurun_out/plain_lm.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_lm_$W.csv python tools/prof_rj.py --workload $W --what lm --reps 2 > gpurun_out/ncu_lm.log 2>&1
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/launches_lm_$W.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(",",""))
    except: continue
    if r[ui]=="ns": v/=1000.0
    elif r[ui]=="ms": v*=1000.0
    k=r[ki].split("(")[0][:40]
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
tot=sum(a[1] for a in agg.values())
for k,a in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print("%-42s n=%4d total %10.1f us  avg %9.1f us  %5.1f%%" % (k,a[0],a[1],a[1]/a[0],100*a[1]/tot))
PY
