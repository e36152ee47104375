#!/bin/bash
# A/B of the BLAS builders on the mesh configs (one GPU): host binned SAH, device binned SAH, device linear BVH.
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -s -k "gpu_built_blas or device_sah_builder" 2>&1 | grep -E "traversal work|rRMSE|passed|failed|Error|assert" | head -20
run() {  # tag config extra-args...
  local tag=$1 cfg=$2; shift 2
  python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/abb_${cfg}_$tag.json 2> gpurun_out/abb_${cfg}_$tag.err; echo "$cfg $tag rc=$?"
}
run host c3 --host-blas
ZB_GPU_BUILDER=1 run dsah c3 --gpu-blas
ZB_GPU_BUILDER=0 run lbvh c3 --gpu-blas
if [ "${WITH_C5:-0}" = 1 ]; then
  run host c5 --spp 16 --host-blas
  ZB_GPU_BUILDER=1 run dsah c5 --spp 16 --gpu-blas
  ZB_GPU_BUILDER=0 run lbvh c5 --spp 16 --gpu-blas
fi
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/abb_c[35]_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms %.2f Mpaths/s %.0f nodes/ray %.2f prims/ray %.2f upload %.1f ms blas %s' % (d['ms_per_step'], d['value'], d['config']['node_visits_per_ray'], d['config']['prim_tests_per_ray'], d['e2e']['parts_ms']['scene_upload'], d['config']['blas']))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json', '.err')).read()[-400:])
PY
