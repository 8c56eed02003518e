#!/bin/bash
# A/B builds of the decoder: scripts/build_variant.sh <name> <nvcc -D flags...>  ->  variants/libtaco_<name>.so
set -e
cd "$(dirname "$0")/.."
N=$1; shift
C=tacotron_prosody_multi_gpu_b200/csrc
mkdir -p variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC "$@" -c $C/decoder3.cu -o variants/decoder3_$N.o
OBJS=$(ls $C/build/*.o | grep -v decoder3.o)
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o variants/libtaco_$N.so $OBJS variants/decoder3_$N.o -lcudart -lcuda
echo variants/libtaco_$N.so
