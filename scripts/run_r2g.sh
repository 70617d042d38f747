#!/bin/bash
# round-2, 4 GPUs: the strips bench at N = 2 and N = 4 (N = 4 is BASELINE configs[3]: one 32768^2 DEM in 4 strips)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python scripts/acc_probe.py 8192 2>&1 | tee gpurun_out/r02_acc_probe_8192c.txt
timeout 500 $TR --nproc-per-node 2 --master-port 29521 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/r02_bench_n2.err | tail -1 > gpurun_out/bench_r02_n2.json; cut -c1-1800 gpurun_out/bench_r02_n2.json; grep -E "Error|error" gpurun_out/r02_bench_n2.err | head -5
timeout 900 $TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --steps 20 --warmup 5 2>gpurun_out/r02_bench_n4.err | tail -1 > gpurun_out/bench_r02_n4.json; cut -c1-3500 gpurun_out/bench_r02_n4.json; grep -E "Error|error" gpurun_out/r02_bench_n4.err | head -5
