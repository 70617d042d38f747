#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.log 2>gpurun_out/r02_bench_n1.err; tail -1 gpurun_out/r02_bench_n1.log > gpurun_out/bench_r02_n1.json; cut -c1-1200 gpurun_out/bench_r02_n1.json; tail -3 gpurun_out/r02_bench_n1.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/bench_r02_reference.json; cut -c1-400 gpurun_out/bench_r02_reference.json
