#!/bin/bash
# times the flow sweeps for every variant library under build/ (git-ignored) and the in-tree one
for f in build/lib_*.so soillib_b200/libsoilb200.so; do
  echo "== $f"; SOILB200_LIB=$PWD/$f python scripts/flow_probe.py 8192 2>&1 | grep -E "^direction|^steepest|^normal"
done
