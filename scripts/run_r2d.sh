#!/bin/bash
# round-2, 2 GPUs: real ranks against the undivided map (peer-mapped / hand-off, halo depths, with and without the
# overlapped halo exchange), then the strips bench at N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
{
timeout 120 $TR --master-port 29511 scripts/strips_check_nccl.py --peers --halo 1 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
timeout 120 $TR --master-port 29512 scripts/strips_check_nccl.py --peers --halo 64 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
timeout 120 $TR --master-port 29513 scripts/strips_check_nccl.py --peers --halo 64 --no-overlap 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
timeout 120 $TR --master-port 29514 scripts/strips_check_nccl.py --halo 1 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5
} > gpurun_out/r02_strips_check_n2.txt 2>&1; cat gpurun_out/r02_strips_check_n2.txt
timeout 400 $TR --master-port 29515 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/r02_bench_n2.err | tail -1 > gpurun_out/bench_r02_n2.json; cut -c1-2500 gpurun_out/bench_r02_n2.json; tail -3 gpurun_out/r02_bench_n2.err
timeout 400 $TR --master-port 29516 bench.py --gpus 2 --steps 20 --warmup 5 --no-overlap --halo 1 2>gpurun_out/r02_bench_n2b.err | tail -1 > gpurun_out/bench_r02_n2_halo1_nooverlap.json; cut -c1-900 gpurun_out/bench_r02_n2_halo1_nooverlap.json
