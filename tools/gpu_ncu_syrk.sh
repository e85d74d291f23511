#!/bin/bash
# ncu --set full of k_syrk96 on a pure diagonal-tile grid (16 cameras) and on the mixed grid (64 cameras): summary + stall samples per source line
mkdir -p gpurun_out
for wl in cfg4shard@16 cfg4shard; do
  tag=$(echo $wl | tr '@' '_')
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_syrk96$" -s 4 -c 1 -o gpurun_out/syrk96_$tag python tools/time_syrk.py $wl > gpurun_out/ncu_syrk96_$tag.log 2>&1
  python tools/ncu_summary.py gpurun_out/syrk96_$tag.ncu-rep gpurun_out/syrk96_${tag}_summary.txt > /dev/null; cat gpurun_out/syrk96_${tag}_summary.txt
  ncu -i gpurun_out/syrk96_$tag.ncu-rep --page source --csv > gpurun_out/syrk96_${tag}_source.csv 2>/dev/null
  python tools/ncu_source_lines.py gpurun_out/syrk96_${tag}_source.csv _Z8k_syrk96 bax_syrk > gpurun_out/syrk96_${tag}_source_lines.txt 2>&1; head -n 30 gpurun_out/syrk96_${tag}_source_lines.txt
  rm -f gpurun_out/syrk96_$tag.ncu-rep
done
