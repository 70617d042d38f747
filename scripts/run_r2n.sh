#!/bin/bash
# round-2, 8 GPUs: the strips bench at N = 4 and N = 8 (weak shape as value, configs[3] 32768^2 as extra key), real-rank check at 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29541 scripts/strips_check_nccl.py --peers --halo 16 --view 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -3 | tee gpurun_out/r02_strips_check_n8_view.txt
for n in 4 8; do
  timeout 900 $TR --nproc-per-node $n --master-port $((29541+n)) bench.py --gpus $n --steps 20 --warmup 5 2>gpurun_out/r02_bench_n$n.err | tail -1 > gpurun_out/bench_r02_n$n.json
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_r02_n$n.json').read())
    print($n, d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['phases_ms']['breakdown']['max_over_ranks'])
    c=d.get('config4_32768')
    if c: print('  C4', c['value'], c['ms_per_step'], 'e2e', c['e2e']['value'], c['phases_ms']['breakdown']['max_over_ranks'])
except Exception as e:
    print('N=$n failed', e)
PY
  grep -E "Error|error" gpurun_out/r02_bench_n$n.err | head -3
done
