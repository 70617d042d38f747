#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
SOILB200_PLAIN_DIV=1 python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
