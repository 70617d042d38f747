#!/bin/bash
# build/lib_<name>.so = the product library with ONE source (VARIANT_SRC, default flow) compiled under extra -D
# flags (tuning builds, git-ignored).  usage: [VARIANT_SRC=grad] scripts/build_variant.sh <name> [-DMACRO=V ...]
set -e
cd "$(dirname "$0")/.."
name=$1; shift
v=${VARIANT_SRC:-flow}
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++20 -fmad=false -Xcompiler -fPIC,-O3,-ffp-contract=off"
mkdir -p build/obj
all="flow accumulate accumulate_tiled transport dense noise multiflow grad albedo path"
objs=""
for f in $all; do
  if [ $f = $v ]; then objs="$objs build/obj/${v}_$name.o"; continue; fi
  [ build/obj/$f.o -nt soillib_b200/csrc/$f.cu ] || nvcc $F -c -o build/obj/$f.o soillib_b200/csrc/$f.cu 2>/dev/null
  objs="$objs build/obj/$f.o"
done
nvcc $F "$@" -c -o build/obj/${v}_$name.o soillib_b200/csrc/$v.cu
nvcc $F -shared -o build/lib_$name.so $objs
echo build/lib_$name.so
