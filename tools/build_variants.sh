#!/bin/bash
# Builds libb200svs variants with different -D flags into build/variants/<name>.so (perf experiments; the
# bench picks one up by copying it over music_generation_b200/lib/libb200svs.so on the GPU box).
# usage: tools/build_variants.sh name1="-DFOO=1" name2="-DFOO=2" ...
set -e
cd "$(dirname "$0")/../music_generation_b200/csrc"
mkdir -p ../../build/variants
for spec in "$@"; do
  name="${spec%%=*}"; flags="${spec#*=}"
  d=../../build/variants/obj_$name; mkdir -p $d
  for f in api stages resblock_pair conv1d_ws conv1d_tc conv1d_f32 attention attention_tc attention_ts pointwise probe song; do
    nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC $flags -c $f.cu -o $d/$f.o &
  done
  wait
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/variants/$name.so $d/*.o -cudart static
  rm -rf $d
  echo built $name
done
