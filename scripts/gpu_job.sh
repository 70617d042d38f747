#!/bin/bash
# The GPU-side jobs of this repository, one name each; run on a B200 box as
#   /usr/local/graft/bin/gpurun [--gpus N] -- 'bash scripts/gpu_job.sh <job> [args]'
# Everything a job leaves behind goes to gpurun_out/ (scratch); what is to be kept is copied to profiles/ by hand.
#   tests                 the whole -m gpu suite
#   smoke                 __graft_entry__.smoke()
#   bench [N]             bench.py at N GPUs (default 1), plus the reference arm at N = 1
#   phases [P ...]        particle / dense phase times of the erosion step at 8192^2 (scripts/particle_sweep.py)
#   micro                 scripts/micro_gather.cu: layouts and chunk sizes against the DRAM request rate
#   acc                   accumulate / accumulate_decay timings at 8192^2 and 1024^2, against the reference's kernels
#   steepest              the three loaders of the D8 / D4 sweep (ring, 1-D bulk rows, 2-D tensor-map boxes)
#   ncu-step              ncu --set full of the step's three kernels            -> gpurun_out/prof_erode.ncu-rep
#   ncu-launches          ncu launch list of a short default bench               -> gpurun_out/launches_bench.csv
#   strips-check N [...]  N real ranks against the undivided map (scripts/strips_check_nccl.py; extra args passed on)
#   halo-sweep N          bench.py --gpus N --no-extra for halo depths 64 / 128 / 256 / 512
#   spawn-probe [N ...]   one GPU as rank 0 of N ranks: what the strip machinery costs the particle phase
#   ncu-acc               ncu --set full of one accumulate call's launches     -> gpurun_out/prof_r02_acc.ncu-rep
set -u
mkdir -p gpurun_out
job=${1:-tests}; shift || true
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
case "$job" in
  tests) timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -12 ;;
  smoke) python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 ;;
  bench)
    n=${1:-1}
    if [ "$n" -eq 1 ]; then
      timeout 900 python bench.py --steps 20 --warmup 5 2>gpurun_out/bench_n1.err | tail -1 > gpurun_out/bench_n1.json
      timeout 300 python bench.py --impl reference --steps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/bench_reference.json
      cut -c1-700 gpurun_out/bench_n1.json; cut -c1-300 gpurun_out/bench_reference.json
    else
      timeout 1200 $TR --nproc-per-node "$n" --master-port $((29500 + n)) bench.py --gpus "$n" --steps 20 --warmup 5 \
        2>gpurun_out/bench_n$n.err | tail -1 > gpurun_out/bench_n$n.json
      cut -c1-700 gpurun_out/bench_n$n.json; grep -E "Error|error" gpurun_out/bench_n$n.err | head -3
    fi ;;
  phases) python scripts/particle_sweep.py 8192 "$@" 2>&1 | tee gpurun_out/particle_sweep.txt ;;
  micro) timeout 300 build/micro_gather 2>&1 | tee gpurun_out/micro_gather.txt ;;
  acc) python scripts/acc_probe.py 8192 --ref 2>&1 | tee gpurun_out/acc_probe_8192.txt; python scripts/acc_probe.py 1024 2>&1 | tee gpurun_out/acc_probe_1024.txt ;;
  steepest) for v in default staged tma; do SOILB200_STEEPEST=$v python scripts/steepest_probe.py 8192; done 2>&1 | tee gpurun_out/steepest_variants.txt ;;
  ncu-step)
    python scripts/particle_sweep.py 8192 32768 > gpurun_out/plain_step.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_fused_step|k_transport_both|k_pre" -s 15 -c 3 -f \
      -o gpurun_out/prof_erode python scripts/particle_sweep.py 8192 32768 > gpurun_out/ncu_step.log 2>&1
    tail -2 gpurun_out/ncu_step.log ;;
  spawn-probe) python scripts/spawn_probe.py "$@" 2>&1 | grep world | tee gpurun_out/spawn_probe.txt ;;
  ncu-acc)
    python scripts/acc_probe.py 8192 > gpurun_out/plain_acc.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_exits|k_tile|k_reduce" -s 20 -c 10 -f \
      -o gpurun_out/prof_r02_acc python scripts/acc_probe.py 8192 > gpurun_out/ncu_acc.log 2>&1
    tail -2 gpurun_out/ncu_acc.log ;;
  ncu-launches)
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/plain_bench.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/ncu_launches.log 2>&1
    tail -2 gpurun_out/ncu_launches.log | cut -c1-200 ;;
  strips-check)
    n=${1:-2}; shift || true
    timeout 200 $TR --nproc-per-node "$n" --master-port $((29600 + n)) scripts/strips_check_nccl.py "$@" 2>&1 | grep -E "NCCL strips|Error|error|assert" | head -5 ;;
  halo-sweep)
    n=${1:-8}
    for halo in 64 128 256 512; do
      timeout 600 $TR --nproc-per-node "$n" --master-port $((29700 + halo % 97)) bench.py --gpus "$n" --steps 20 --warmup 5 --no-extra --halo $halo \
        2>gpurun_out/bench_n${n}_h$halo.err | tail -1 > gpurun_out/bench_n${n}_h$halo.json
      python -c "import json; d=json.load(open('gpurun_out/bench_n${n}_h$halo.json')); print('halo $halo', d['value'], d['ms_per_step'], d['phases_ms']['breakdown']['max_over_ranks'])"
    done ;;
  *) echo "unknown job $job"; exit 2 ;;
esac
