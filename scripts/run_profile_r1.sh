#!/bin/bash
# round-1 profile pass (1 GPU): launch list of the default bench command, then full captures of the hot kernels
mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 80 --csv --log-file gpurun_out/launches_r01.csv python bench.py --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -1 gpurun_out/plain_bench.log > gpurun_out/bench_r01_n1.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both" -s 8 -c 2 -o gpurun_out/prof_r01_erode python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
python scripts/flow_probe.py 8192 --ref > gpurun_out/flow_probe_8192.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_sweep<|k_tile|k_reduce|k_creep|k_normal" -c 14 -o gpurun_out/prof_r01_flow python scripts/flow_probe.py 8192 > gpurun_out/ncu3.log 2>&1
cat gpurun_out/flow_probe_8192.txt
python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_r01_reference.json
cut -c1-300 gpurun_out/bench_r01_n1.json; cut -c1-200 gpurun_out/bench_r01_reference.json
