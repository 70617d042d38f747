#!/bin/bash
# build/lib_<name>.so = the product library with flow.cu compiled under extra -D flags (tuning builds, git-ignored)
# usage: scripts/build_variant.sh <name> [-DMACRO=V ...]
set -e
cd "$(dirname "$0")/.."
name=$1; shift
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++20 -fmad=false -Xcompiler -fPIC,-O3,-ffp-contract=off"
mkdir -p build/obj
for f in accumulate accumulate_tiled transport dense noise multiflow grad albedo path; do
  [ build/obj/$f.o -nt soillib_b200/csrc/$f.cu ] || nvcc $F -c -o build/obj/$f.o soillib_b200/csrc/$f.cu 2>/dev/null
done
nvcc $F "$@" -c -o build/obj/flow_$name.o soillib_b200/csrc/flow.cu
nvcc $F -shared -o build/lib_$name.so build/obj/flow_$name.o build/obj/{accumulate,accumulate_tiled,transport,dense,noise,multiflow,grad,albedo,path}.o
echo build/lib_$name.so
