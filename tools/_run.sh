cd /root/repo
run() { # lib scene spp
  ZB_LIB=$1 timeout 600 python tools/tune.py --scene $2 --spp $3 --reps 4 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$1 $2 $3', round(d['ms'],3), 'ms', d['mean'])"
}
for lib in raytracer_project_b200/libzenith_b200.so raytracer_project_b200/_build/lib_p7.so; do
  run $PWD/$lib weekend 64
  run $PWD/$lib cornell 32
  run $PWD/$lib mesh 16
done
ZB_LIB=$PWD/raytracer_project_b200/_build/lib_p7.so timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
