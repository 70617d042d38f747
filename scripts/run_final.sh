#!/bin/bash
# end-of-round pass on one GPU: full GPU test suite, default bench (both arms), one full capture of the dense kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
python bench.py > gpurun_out/plain_bench.log 2>&1; tail -1 gpurun_out/plain_bench.log > gpurun_out/bench_r01_n1.json
python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_r01_reference.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both|k_pre" -s 9 -c 3 -o gpurun_out/prof_r01_erode python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
cut -c1-1700 gpurun_out/bench_r01_n1.json; echo; cut -c1-300 gpurun_out/bench_r01_reference.json
