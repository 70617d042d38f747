#!/bin/bash
# round-2, 8 GPUs: halo depth against the overlapped exchange (weak shape only), then the three steepest variants on GPU 0
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for halo in 64 128 256; do
  timeout 600 $TR --nproc-per-node 8 --master-port $((29560+halo)) bench.py --gpus 8 --steps 20 --warmup 5 --no-extra --halo $halo 2>gpurun_out/r02_bench_n8_h$halo.err | tail -1 > gpurun_out/bench_r02_n8_h$halo.json
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_r02_n8_h$halo.json').read())
    print('halo $halo', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['phases_ms']['breakdown']['max_over_ranks'])
except Exception as e:
    print('halo $halo failed', e)
PY
done
for v in default staged tma; do SOILB200_STEEPEST=$v python scripts/steepest_probe.py 8192; done 2>&1 | tee gpurun_out/r02_steepest_variants.txt
timeout 600 python -m pytest tests/test_gpu_flow.py -x -q -m gpu -k "staged" 2>&1 | tail -3
