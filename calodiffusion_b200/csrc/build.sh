#!/bin/bash
# Build libcalodiffusion_b200.so for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
OUT=../libcalodiffusion_b200.so
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
nvcc $FLAGS -c cd_kernels.cu -o cd_kernels.o &
nvcc $FLAGS -c cd_conv_mma.cu -o cd_conv_mma.o &
nvcc $FLAGS -c cd_engine.cu -o cd_engine.o &
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xlinker --no-undefined -o $OUT cd_kernels.o cd_conv_mma.o cd_engine.o -lcudart
echo built $OUT
