#!/bin/bash
# Build libzenith_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
#   ZB_EXPERIMENTS=1  also compiles the rejected kernel forms (wavefront.cu, k_render_pqr) for A/B runs
#   ZB_TAG / ZB_OUT   build directory tag / output path (side-by-side variants for tuning)
set -euo pipefail
cd "$(dirname "$0")/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
EXP=""
if [ "${ZB_EXPERIMENTS:-0}" != "0" ]; then EXP="-DZB_EXPERIMENTS"; fi
FLAGS="-O3 -std=c++20 -lineinfo $EXP -Xcompiler -fPIC,-Wall,-Wno-unused-function ${ZB_NVCC_EXTRA:-}"
OUT=${ZB_OUT:-../libzenith_b200.so}
B=../_build/${ZB_TAG:-main}
mkdir -p $B
CU="kernels abi abi_render finish lbvh wireframe"
if [ -n "$EXP" ]; then CU="$CU wavefront"; fi
for f in $(ls *.cu | sed 's/\.cu$//'); do
  case " $CU " in *" $f "*) ;; *) if [ "$f" != "wavefront" ]; then CU="$CU $f"; fi ;; esac
done
OBJS=""
for f in $CU; do
  rm -f $B/$f.o
  $NVCC $ARCH $FLAGS -c $f.cu -o $B/$f.o &
  OBJS="$OBJS $B/$f.o"
done
g++ -O2 -std=c++20 -fPIC -Wall $EXP -c host_glue.cpp -o $B/host_glue.o &
wait
$NVCC $ARCH -shared -o $OUT $OBJS $B/host_glue.o -cudart static -lz -lrt -lpthread
# SHA-1 of every kernel's SASS (ties ncu instruction counts under profiles/ to the code that runs; tools/sass_hash.py)
if [ "$OUT" = "../libzenith_b200.so" ]; then python3 ../../tools/sass_hash.py $OUT ../_build/sass_hashes.json || true; fi
echo "built $OUT"
