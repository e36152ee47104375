cd /root/repo
run() {
  timeout 600 python tools/tune.py --scene mesh --spp 16 --reps 3 --stats 2>/dev/null | tail -1 | python -c "
import sys,json,os
d=json.loads(sys.stdin.read()); print('ci', os.environ.get('ZB_BLAS_CI'), 'leaf', os.environ.get('ZB_BLAS_LEAF'), round(d['ms'],3), 'ms nodes/ray', round(d['nodes_per_ray'],2), 'prims/ray', round(d['prims_per_ray'],2), 'MB', d['scene_mb'])"
}
run
ZB_BLAS_CI=0.5 run
ZB_BLAS_CI=0.7 run
ZB_BLAS_CI=1.5 run
ZB_BLAS_CI=0.5 ZB_BLAS_LEAF=8 run
ZB_BLAS_CI=0.7 ZB_BLAS_LEAF=8 run
ZB_BLAS_LEAF=2 run
ZB_BLAS_LEAF=8 run
