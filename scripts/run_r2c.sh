#!/bin/bash
# round-2: the whole -m gpu suite, then the default bench line (1 GPU) and the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r02_gpu_tests.txt; cat gpurun_out/r02_gpu_tests.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.log 2>gpurun_out/r02_bench_n1.err; tail -1 gpurun_out/r02_bench_n1.log > gpurun_out/bench_r02_n1.json; cut -c1-3000 gpurun_out/bench_r02_n1.json; tail -5 gpurun_out/r02_bench_n1.err
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 | cut -c1-600
