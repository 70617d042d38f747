#!/bin/bash
# ncu --set full of the flow sweeps of the tuning builds given as arguments, summarised on the box
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for lib in "$@"; do
  tag=$(basename $lib .so)
  export SOILB200_LIB=$PWD/$lib
  timeout 120 python scripts/flow_once.py 8192 > gpurun_out/once.log 2>&1 || { tail -5 gpurun_out/once.log; exit 1; }
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_steepest" -c 1 -f -o gpurun_out/prof_$tag python scripts/flow_once.py 8192 > gpurun_out/ncu.log 2>&1
  tail -1 gpurun_out/ncu.log
  python scripts/summarize_ncu.py gpurun_out/prof_$tag.ncu-rep gpurun_out/prof_$tag.txt
  ncu -i gpurun_out/prof_$tag.ncu-rep --page raw --csv | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
for r in rows[2:]:
    for k,v in zip(h,r):
        if ('pipe' in k and 'pct' in k) or 'achieved_occupancy' in k or 'warps_eligible' in k or 'issue_active' in k:
            print(k,'=',v)
" >> gpurun_out/prof_$tag.txt
done
