#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_erosion.py tests/test_gpu_reference.py tests/test_gpu_strips.py tests/test_gpu_baseline_sizes.py -x -q -m gpu 2>&1 | tail -3
python scripts/particle_sweep.py 8192 8192 32768 2>&1 | tail -2
