#!/bin/bash
# validation after the gradient / negslope feeder: full GPU test suite, smoke, flow probe
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -1
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 200 python scripts/flow_probe.py 8192 > gpurun_out/flow_probe_8192b.txt 2>&1; cat gpurun_out/flow_probe_8192b.txt
