#!/bin/bash
# full GPU test suite on the in-tree library, then the tuning builds of the normal sweep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/gpu_tests.txt
timeout 200 python scripts/variant_probe.py 8192 > gpurun_out/variants.txt 2>&1; tail -6 gpurun_out/variants.txt | cut -c90-200
