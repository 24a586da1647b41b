#!/bin/bash
# Build variants of libb200ba.so that differ in -D flags of ba_kernels.cu and print the script to time them on the GPU box:
#   tools/ab_variants.sh name1:"-DA=1" name2:"-DB=2 -DC=3" ...     ->  sfm-python_b200/variants/lib_<name>.so
set -e
cd "$(dirname "$0")/../sfm-python_b200/csrc"
mkdir -p ../variants
make -j4 > /dev/null
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v $flags \
      -c ba_kernels.cu -o ../variants/ba_kernels_$name.o 2> ../variants/$name.ptxas.log
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/lib_$name.so ba_setup.o ../variants/ba_kernels_$name.o ba_solve.o ba_engine.o -lcudart -ldl
  grep -A3 "k_lin_flow" ../variants/$name.ptxas.log | grep -E "Used|spill" | tr '\n' ' '; echo " <- $name"
done
