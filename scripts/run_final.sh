#!/bin/bash
# end-of-round pass on one GPU: full GPU test suite, smoke, default bench (both arms), flow probes
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/plain_bench.log 2>&1; tail -1 gpurun_out/plain_bench.log > gpurun_out/bench_r01_n1.json
python bench.py --impl reference --steps 10 --warmup 3 2>&1 | tail -1 > gpurun_out/bench_r01_reference.json
python scripts/flow_probe.py 8192 --ref > gpurun_out/flow_probe_8192.txt 2>&1
grep -iE "gradient|negslope|laplacian|slope|blur" gpurun_out/flow_probe_8192.txt > gpurun_out/grad_probe_8192.txt
python scripts/flow_probe.py 1024 > gpurun_out/flow_probe_1024.txt 2>&1
cut -c1-1800 gpurun_out/bench_r01_n1.json; echo; cut -c1-260 gpurun_out/bench_r01_reference.json; echo; cat gpurun_out/flow_probe_8192.txt
