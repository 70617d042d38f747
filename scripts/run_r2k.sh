#!/bin/bash
# round-2 profile pass: full ncu captures of the step's three kernels (view mode) at 8192^2
mkdir -p gpurun_out
python scripts/particle_sweep.py 8192 32768 > gpurun_out/plain_r2k.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both|k_pre" -s 15 -c 3 -f -o gpurun_out/prof_r02_erode python scripts/particle_sweep.py 8192 32768 > gpurun_out/ncu_r2k.log 2>&1; tail -2 gpurun_out/ncu_r2k.log
