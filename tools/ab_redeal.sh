#!/bin/bash
# block-synchronous segments (1) and octant re-dealing of rays (2) in k_render_pq, experiments build.
set -u
export ZB_LIB=$PWD/raytracer_project_b200/libzenith_b200_exp.so ZB_BENCH_CACHE=/tmp/zb_cache
for m in 0 1 2; do
  for c in headline c3 c2; do
    ZB_PQ_REDEAL=$m python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('redeal=$m', d['config']['workload'], 'ms %.2f Mpaths/s %.0f' % (d['ms_per_step'], d['value']), d['mean_radiance'][0])"
  done
done
