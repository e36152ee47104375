#!/bin/bash
# One GPU: every BASELINE config through bench.py, then ncu captures of the mesh kernel on C3 and (at 16 spp) C5.
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
for c in c1 c2 c3 c4; do
  python bench.py --config $c --steps 3 --warmup 3 ${NOCPU:-} > gpurun_out/cfg_$c.json 2> gpurun_out/cfg_$c.err; echo "$c rc=$?"
done
python bench.py --config c5 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/cfg_c5.json 2> gpurun_out/cfg_c5.err; echo "c5 rc=$?"
if [ "${SKIP_NCU:-0}" = 1 ]; then exit 0; fi
python bench.py --config c3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_pq -s 1 -c 1 -f -o gpurun_out/r02_pq_mesh0_c3 python bench.py --config c3 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
python bench.py --config c5 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_c5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_pq -s 1 -c 1 -f -o gpurun_out/r02_pq_mesh0_c5 python bench.py --config c5 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_c5.log 2>&1; echo "ncu c5 rc=$?"
