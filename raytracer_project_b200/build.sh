#!/bin/bash
# Build libzenith_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++20 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function ${ZB_NVCC_EXTRA:-}"
OUT=${ZB_OUT:-../libzenith_b200.so}
B=../_build/${ZB_TAG:-main}
mkdir -p $B
$NVCC $ARCH $FLAGS -c kernels.cu -o $B/kernels.o &
$NVCC $ARCH $FLAGS -c abi.cu -o $B/abi.o &
$NVCC $ARCH $FLAGS -c abi_render.cu -o $B/abi_render.o &
$NVCC $ARCH $FLAGS -c finish.cu -o $B/finish.o &
$NVCC $ARCH $FLAGS -c lbvh.cu -o $B/lbvh.o &
$NVCC $ARCH $FLAGS -c wireframe.cu -o $B/wireframe.o &
$NVCC $ARCH $FLAGS -Xptxas -v -c wavefront.cu -o $B/wavefront.o 2> $B/wavefront.ptxas &
g++ -O2 -std=c++20 -fPIC -Wall -c host_glue.cpp -o $B/host_glue.o &
wait
$NVCC $ARCH -shared -o $OUT $B/kernels.o $B/abi.o $B/abi_render.o $B/finish.o $B/lbvh.o $B/wavefront.o $B/wireframe.o $B/host_glue.o -cudart static -lz
echo "built $OUT"
