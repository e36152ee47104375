#!/bin/bash
# Device SAH builder: cost of a triangle test (percent of a node visit) and leaf size against the C3 / C5 frame time.
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
for leaf in 4 8 2; do
  for pct in 100 60 150; do
    for cfg in c3 c5; do
      extra=""; [ $cfg = c5 ] && extra="--spp 16"
      ZB_DEVICE_BLAS_LEAF=$leaf ZB_BLAS_ISECT_PCT=$pct python bench.py --config $cfg $extra --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/sahp_${cfg}_l${leaf}_p${pct}.json 2> gpurun_out/sahp_${cfg}_l${leaf}_p${pct}.err
    done
  done
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/sahp_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms %.2f nodes/ray %.2f prims/ray %.2f' % (d['ms_per_step'], d['config']['node_visits_per_ray'], d['config']['prim_tests_per_ray']))
    except Exception as e:
        print(f, 'ERR', e)
PY
