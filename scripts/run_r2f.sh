#!/bin/bash
# round-2: whole GPU suite after the accumulate / map_t / examples changes, accumulate timings, smoke
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -12
python scripts/acc_probe.py 8192 2>&1 | tee gpurun_out/r02_acc_probe_8192b.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_exits|k_tile|k_reduce" -s 40 -c 10 --csv --log-file gpurun_out/r02_acc_launches.csv python scripts/acc_probe.py 8192 > /dev/null 2>&1; tail -10 gpurun_out/r02_acc_launches.csv | cut -d, -f5,15- | cut -c1-60,140-
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
