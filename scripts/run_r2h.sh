#!/bin/bash
# round-2: packed particle view + k_dense4: whole GPU suite, particle/dense phase times with and without the view
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -12
python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
SOILB200_NO_VIEW=1 python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
python scripts/particle_sweep.py 8192 8192 131072 2>&1 | tail -2
