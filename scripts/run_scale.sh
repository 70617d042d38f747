#!/bin/bash
# weak-scaling pass on one box: `run_scale.sh "1 2 4 8"` runs the strips bench (peer-mapped) at each N listed, then the
# replica-parallel multi-flow at the largest N
mkdir -p gpurun_out
port=29500
last=1
for n in $1; do
  port=$((port+1)); last=$n
  if [ $n -eq 1 ]; then
    timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/scale_err_$n.log | tail -1 > gpurun_out/scale_n$n.json
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/scale_err_$n.log | tail -1 > gpurun_out/scale_n$n.json
  fi
  python -c "
import json,sys
d=json.loads(open('gpurun_out/scale_n$n.json').read()); print($n, d['value'], d['ms_per_step'], d.get('phases_ms'), d['roofline']['frac'], d['e2e']['value'])" || tail -5 gpurun_out/scale_err_$n.log
done
if [ $last -gt 1 ]; then
  port=$((port+1))
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $last --master-addr 127.0.0.1 --master-port $port scripts/multiflow_replicas.py 8192 16 2>&1 | grep -E "^multiflow|Error|error" | head -5
fi
