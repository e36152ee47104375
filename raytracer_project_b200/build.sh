#!/bin/bash
# Build libzenith_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++20 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function ${ZB_NVCC_EXTRA:-}"
mkdir -p ../_build
$NVCC $ARCH $FLAGS -c kernels.cu -o ../_build/kernels.o &
$NVCC $ARCH $FLAGS -c abi.cu -o ../_build/abi.o &
$NVCC $ARCH $FLAGS -c abi_render.cu -o ../_build/abi_render.o &
g++ -O2 -std=c++20 -fPIC -Wall -c host_glue.cpp -o ../_build/host_glue.o &
wait
$NVCC $ARCH -shared -o ../libzenith_b200.so ../_build/kernels.o ../_build/abi.o ../_build/abi_render.o ../_build/host_glue.o -cudart static
echo "built $(cd .. && pwd)/libzenith_b200.so"
