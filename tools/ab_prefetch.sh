#!/bin/bash
# A/B: prefetch of the deferred child in mesh traversals (ZB_PREFETCH_FAR), two builds of the library side by side.
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
for v in pf nopf; do
  lib=$PWD/raytracer_project_b200/libzenith_b200.so; [ $v = nopf ] && lib=$PWD/raytracer_project_b200/libzenith_b200_nopf.so
  ZB_LIB=$lib python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/abp_c3_$v.json 2> gpurun_out/abp_c3_$v.err; echo "c3 $v rc=$?"
  :
  ZB_LIB=$lib python bench.py --config c5 --spp 16 --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/abp_c5_$v.json 2> gpurun_out/abp_c5_$v.err; echo "c5 $v rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/abp_c*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms %.2f Mpaths/s %.0f kernel %s' % (d['ms_per_step'], d['value'], d['config']['kernel']))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json', '.err')).read()[-300:])
PY
