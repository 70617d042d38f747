#!/bin/bash
# flow-sweep tuning builds against the in-tree library (bit compare + CUDA-event timings)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 220 python scripts/variant_probe.py 8192 "$@" > gpurun_out/variants.txt 2>&1; tail -16 gpurun_out/variants.txt | cut -c1-120
