#!/bin/bash
# Build a variant of libb200attn.so with extra nvcc flags (timing ablations, A/B):
#   benchmarks/_ab/build_variant.sh <name> [-DB200_ABLATE_FILL=1 ...]  ->  benchmarks/_ab/lib_<name>.so
# (only extend.cu is recompiled with the flags unless ALL=1 is set in the environment)
set -e
name=$1; shift
out=benchmarks/_ab/obj_$name; mkdir -p $out
for f in abi kv_store decode extend merge_state oproj_allreduce; do
  src=sglang_b200/csrc/$f.cu
  if [ $f = extend ] || [ -n "$ALL" ] || [ ! -f sglang_b200/lib/$f.cu.o ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I include "$@" -c $src -o $out/$f.o &
  else
    cp sglang_b200/lib/$f.cu.o $out/$f.o
  fi
done
wait
nvcc -shared -o benchmarks/_ab/lib_$name.so $out/*.o
echo built benchmarks/_ab/lib_$name.so
