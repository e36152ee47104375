#!/bin/bash
# One 8-GPU box, back to back (gpurun --gpus 8): headline at N = 1, 2, 4, 8; config C1 (10 spp: tile stripes) at N = 8;
# config C5 (10 M triangles, 4K x 1024 spp) at N = 8, 4, 2 with sample slices and at N = 8 with tile stripes.
# Results: gpurun_out/scale_*.json (one bench.py JSON line each).
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
run() {  # name, n, bench args...
  local name=$1 n=$2; shift 2
  if [ "$n" = 1 ]; then python bench.py --gpus 1 --no-cpu-baseline "$@" > gpurun_out/scale_$name.json 2> gpurun_out/scale_$name.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n "$@" > gpurun_out/scale_$name.json 2> gpurun_out/scale_$name.err; fi
  echo "$name rc=$? $(tail -c 200 gpurun_out/scale_$name.json | tr -d '\n' | tail -c 120)"
}
nvidia-smi topo -m > gpurun_out/scale_topo.txt 2>&1
timeout 300 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -2
for n in 8 4 2 1; do run headline_n$n $n --steps 5 --warmup 3; done
run c1_n8 8 --config c1 --steps 5 --warmup 3
run c1_n1 1 --config c1 --steps 5 --warmup 3
if [ "${C5_ONLY_N8:-0}" = 1 ]; then run c5_n8 8 --config c5 --steps 2 --warmup 1; fi
if [ "${SKIP_C5:-0}" = 0 ]; then
  run c5_n8 8 --config c5 --steps 2 --warmup 1
  run c5_n8_tiles 8 --config c5 --steps 2 --warmup 1 --partition tiles
  run c5_n4 4 --config c5 --steps 2 --warmup 1
  run c5_n2 2 --config c5 --steps 1 --warmup 1
fi
