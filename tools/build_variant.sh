#!/bin/bash
# usage: tools/build_variant.sh <name> <extra nvcc flags...>   -> epievo_b200/variants/lib_<name>.so
# (A/B testing of kernel build options on the GPU box: EPIEVO_B200_LIB=<that file> python bench.py ...)
set -e
cd "$(dirname "$0")/../epievo_b200"
name=$1; shift
mkdir -p variants/obj_$name
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fno-fast-math --fmad=true"
objs=""
for f in csrc/*.cu csrc/*.cpp; do
  o=variants/obj_$name/$(basename ${f%.*}).o
  nvcc $FLAGS "$@" -c $f -o $o &
  objs="$objs $o"
done
wait
nvcc -shared -o variants/lib_$name.so $objs -lcudart
echo variants/lib_$name.so
