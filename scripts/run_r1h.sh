#!/bin/bash
# round-1 closing pass on one GPU after the flow-sweep rework: tests, smoke, default bench, flow probes, ncu of the sweeps
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -1
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py > gpurun_out/plain_bench.log 2>&1; tail -1 gpurun_out/plain_bench.log > gpurun_out/bench_r01_n1.json
timeout 200 python scripts/flow_probe.py 8192 --ref > gpurun_out/flow_probe_8192.txt 2>&1
timeout 100 python scripts/flow_probe.py 1024 > gpurun_out/flow_probe_1024.txt 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:"k_steepest|k_normal_sweep" -c 3 -f -o gpurun_out/prof_r01_flow2 python scripts/flow_once.py 8192 > gpurun_out/ncu3.log 2>&1
python scripts/summarize_ncu.py gpurun_out/prof_r01_flow2.ncu-rep gpurun_out/prof_r01_flow2.txt
cut -c1-1500 gpurun_out/bench_r01_n1.json; echo; head -12 gpurun_out/flow_probe_8192.txt
