#!/bin/bash
# round-2: new tiled accumulate (pointer-jumping exit totals, tiled decay): tests, timings, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flow.py tests/test_gpu_reference.py tests/test_gpu_baseline_sizes.py tests/test_zz_no_alloc.py tests/test_gpu_strips.py -x -q -m gpu 2>&1 | tail -8
python scripts/acc_probe.py 8192 --ref 2>&1 | tee gpurun_out/r02_acc_probe_8192.txt
python scripts/acc_probe.py 1024 2>&1 | tee gpurun_out/r02_acc_probe_1024.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_exits|k_tile|k_reduce" -c 40 --csv --log-file gpurun_out/r02_acc_launches.csv python scripts/acc_probe.py 8192 > /dev/null 2>&1; tail -22 gpurun_out/r02_acc_launches.csv | cut -d, -f5,15- | head -30
