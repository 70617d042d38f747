#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | grep -E "^E   |FAILED|passed|failed|Mismatch|Max abs|Max rel" | head -40
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | tee gpurun_out/bench_ours_8192.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases_ms'], d['roofline']['achieved'], d['roofline']['hot_cells_per_step'], d['e2e']['value'], d['flow'])"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --size 4096 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases_ms'], d['roofline']['achieved'], d['roofline']['hot_cells_per_step'])"
