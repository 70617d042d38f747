#!/bin/bash
# round-1 profile pass: launch list of the default bench command, then full captures of the hot kernels
mkdir -p gpurun_out
python scripts/flow_probe.py 8192 --ref 2>&1 | tee gpurun_out/flow_probe_8192.txt
python scripts/flow_probe.py 1024 --ref 2>&1 | tee gpurun_out/flow_probe_1024.txt
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 60 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -1 gpurun_out/plain_bench.log | cut -c1-400
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both" -s 8 -c 2 -o gpurun_out/prof_r01_erode python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
python scripts/flow_probe.py 4096 > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_sweep|k_walk|k_indegree|k_creep|k_normal" -c 12 -o gpurun_out/prof_r01_flow python scripts/flow_probe.py 4096 > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
