#!/bin/bash
# A/B of two builds of libb200attn on the same box: $1 = old lib, default lib = new
OLD=${1:-benchmarks/_ab/libb200attn_old.so}
for rep in 1 2; do
  for lib in new old; do
    if [ $lib = old ]; then export B200ATTN_LIB=$OLD; else unset B200ATTN_LIB; fi
    echo "== $lib rep $rep"
    python benchmarks/extend_one.py 8 16384 1024 32 8 5 | tail -1
    python benchmarks/extend_one.py 8 16384 1024 32 8 5 1 | tail -1
    python benchmarks/extend_one.py 8 0 4096 32 8 5 | tail -1
    python benchmarks/decode_one.py 256 8192 32 8 1 1 10 | tail -1
    python benchmarks/decode_one.py 64 4096 32 8 0 1 10 | tail -1
  done
done
