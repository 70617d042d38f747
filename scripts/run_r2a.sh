#!/bin/bash
# round-2 first measurements: particle-count sweep, gather micro-benchmark, full ncu capture of the particle kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python scripts/particle_sweep.py 8192 > gpurun_out/r02_particle_sweep.txt 2>&1; cat gpurun_out/r02_particle_sweep.txt
timeout 300 build/micro_gather > gpurun_out/r02_micro_gather.txt 2>&1; cat gpurun_out/r02_micro_gather.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_transport_both" -s 5 -c 1 -f -o gpurun_out/prof_r02_transport python scripts/particle_sweep.py 8192 32768 > gpurun_out/ncu_r2a.log 2>&1; tail -3 gpurun_out/ncu_r2a.log
