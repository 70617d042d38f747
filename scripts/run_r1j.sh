#!/bin/bash
# validation of the creep sweep: full GPU test suite, then the flow probe (mass_creep line)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -12
timeout 200 python scripts/flow_probe.py 8192 2>&1 | grep -E "mass_creep|direction|gradient"
