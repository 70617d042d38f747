#!/bin/bash
# round-1 profile pass (1 GPU): default bench, its launch list, full captures of the hot kernels, flow probe, reference arm
mkdir -p gpurun_out
python bench.py > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 120 --csv --log-file gpurun_out/launches_r01.csv python bench.py --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -1 gpurun_out/plain_bench.log > gpurun_out/bench_r01_n1.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both|k_pre" -s 9 -c 3 -o gpurun_out/prof_r01_erode python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
python scripts/flow_probe.py 8192 --ref > gpurun_out/flow_probe_8192.txt 2>&1
python scripts/flow_once.py 8192 > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_steepest|k_normal_sweep|k_creep|k_sweep_gibbs|k_tile|k_reduce" -s 10 -c 10 -o gpurun_out/prof_r01_flow python scripts/flow_once.py 8192 > gpurun_out/ncu3.log 2>&1
cat gpurun_out/flow_probe_8192.txt
python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_r01_reference.json
cut -c1-400 gpurun_out/bench_r01_n1.json; cut -c1-200 gpurun_out/bench_r01_reference.json
