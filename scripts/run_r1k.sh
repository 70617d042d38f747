#!/bin/bash
# ncu --set full of k_creep_sweep at 8192^2 (the plain run of the same script exited 0 in scripts/run_r1j.sh's probe)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 45 ncu --set full --clock-control none --import-source on -k regex:"k_creep_sweep" -c 1 -f -o gpurun_out/prof_r01_creep python scripts/flow_once.py 8192 > gpurun_out/ncu4.log 2>&1
tail -1 gpurun_out/ncu4.log
python scripts/summarize_ncu.py gpurun_out/prof_r01_creep.ncu-rep gpurun_out/prof_r01_creep.txt && cat gpurun_out/prof_r01_creep.txt
