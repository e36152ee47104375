#!/bin/bash
# A/B of the 8-wide compressed tree against the binary tree on the mesh configs (one GPU).
set -u
mkdir -p gpurun_out
export ZB_BENCH_CACHE=/tmp/zb_cache
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -s -k "wide_tree or primary_ids" 2>&1 | grep -E "face ids|passed|failed|Error|assert" | head -20
for w in 1 0; do
  for b in ${BLOCKS:-0}; do
    ZB_WIDE_BUILD=1 ZB_WIDE=$w ZB_PQ_BLOCKS=$b python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --host-blas > gpurun_out/ab_c3_w${w}_b$b.json 2> gpurun_out/ab_c3_w${w}_b$b.err; echo "c3 wide=$w blocks=$b rc=$?"
    ZB_WIDE_BUILD=1 ZB_WIDE=$w ZB_PQ_BLOCKS=$b python bench.py --config c5 --spp 16 --steps 3 --warmup 1 --no-cpu-baseline --host-blas > gpurun_out/ab_c5_w${w}_b$b.json 2> gpurun_out/ab_c5_w${w}_b$b.err; echo "c5@16spp wide=$w blocks=$b rc=$?"
  done
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/ab_c[35]_w*_b*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms %.2f Mpaths/s %.0f nodes/ray %.2f prims/ray %.2f kernel %s' % (d['ms_per_step'], d['value'], d['config']['node_visits_per_ray'], d['config']['prim_tests_per_ray'], d['config']['kernel']))
    except Exception as e:
        print(f, 'ERR', e)
PY
