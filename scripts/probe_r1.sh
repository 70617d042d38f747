#!/bin/bash
# round-1 probe: which part of the fused dense kernel is slow; details of failing parity tests
mkdir -p gpurun_out
for opt in "" "--no-albedo" "--particles 32" "--no-albedo --particles 32"; do
  echo "== bench $opt"
  timeout 300 python bench.py --steps 5 --warmup 3 --size 4096 --no-cpu-baseline $opt 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], d['phases_ms'], d['roofline']['achieved'], d['roofline']['hot_cells_per_step'])
    elif 'Error' in l or 'error' in l: print(l.strip())
"
done
timeout 600 python -m pytest tests/test_gpu_erosion.py -m gpu -q -k "debris or fused" 2>&1 | grep -E "Mismatch|Max abs|Max rel|^E  |FAILED|passed|failed|ACTUAL|DESIRED|^tests/" | head -80
