#!/bin/bash
# Build libcalodiffusion_b200.so for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
OUT=../libcalodiffusion_b200.so
FLAGS="$CD_EXTRA -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
pids=()
for f in cd_kernels cd_attention cd_attention_fused cd_postprocess cd_conv_mma cd_conv_zstack cd_conv_gather cd_backward cd_engine; do
  nvcc $FLAGS -c $f.cu -o $f.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xlinker --no-undefined -o $OUT cd_kernels.o cd_attention.o cd_attention_fused.o cd_postprocess.o cd_conv_mma.o cd_conv_zstack.o cd_conv_gather.o cd_backward.o cd_engine.o -lcudart
echo built $OUT
