#!/bin/bash
# round-2, 2 GPUs: strips with the packed view - one-GPU logical strips tests, real ranks vs the undivided map, bench N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_strips.py tests/test_gpu_erosion.py -x -q -m gpu 2>&1 | tail -5
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
{
timeout 120 $TR --master-port 29531 scripts/strips_check_nccl.py --peers --halo 64 --view 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
timeout 120 $TR --master-port 29532 scripts/strips_check_nccl.py --peers --halo 2 --view --no-overlap 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
timeout 120 $TR --master-port 29533 scripts/strips_check_nccl.py --halo 4 --view 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
} > gpurun_out/r02_strips_check_n2_view.txt 2>&1; cat gpurun_out/r02_strips_check_n2_view.txt
timeout 400 $TR --master-port 29535 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/r02_bench_n2.err | tail -1 > gpurun_out/bench_r02_n2.json; cut -c1-1500 gpurun_out/bench_r02_n2.json; grep -E "Error|error" gpurun_out/r02_bench_n2.err | head -5
