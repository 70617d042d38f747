#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flow.py tests/test_gpu_reference.py tests/test_gpu_baseline_sizes.py tests/test_zz_no_alloc.py tests/test_gpu_examples.py -x -q -m gpu 2>&1 | tail -4
echo "--- pointer-jumping sums + inflow push (new)"; python scripts/acc_probe.py 8192 --ref 2>&1 | grep -v REF
echo "--- ordered walk (SOILB200_ACC=walk)"; SOILB200_ACC=walk python scripts/acc_probe.py 8192 2>&1
python scripts/acc_probe.py 1024 2>&1 | head -2
