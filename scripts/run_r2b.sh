#!/bin/bash
# round-2: gather micro-benchmark (Morton layout, chunk curve), parity tests of the fused step incl. the pre-scratch path, particle phase after the division fix
mkdir -p gpurun_out
timeout 300 build/micro_gather > gpurun_out/r02_micro_gather2.txt 2>&1; grep -E "T=   65536|T=  262144|chunks" gpurun_out/r02_micro_gather2.txt
python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_erosion.py tests/test_gpu_reference.py tests/test_gpu_strips.py -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -40 > gpurun_out/r02_tests_b.txt; cat gpurun_out/r02_tests_b.txt
