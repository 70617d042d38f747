#!/bin/bash
echo "view + creep-only pre (mode 2, default)"; python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
echo "view + full pre record (mode 1)"; SOILB200_PRE_MODE=1 python scripts/particle_sweep.py 8192 32768 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_erosion.py tests/test_gpu_reference.py tests/test_gpu_strips.py -x -q -m gpu 2>&1 | tail -3
