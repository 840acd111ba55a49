#!/bin/bash
# build an A/B variant of the library: tools/build_variant.sh NAME file.cu "-DFLAG=1 ..."  ->  cfd_simulations_b200/build/variants/NAME.so
# (the other objects are the current ones of cfd_simulations_b200/build/; run the normal build first)
set -e
name=$1; src=$2; flags=$3
b=cfd_simulations_b200/build; mkdir -p $b/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC $flags -c -o $b/variants/$name.o cfd_simulations_b200/csrc/$src
objs=$(ls $b/*.o | grep -v "/${src%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $b/variants/$name.so $objs $b/variants/$name.o
echo $b/variants/$name.so
