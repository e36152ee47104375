cd /root/repo
timeout 1400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for sc in "weekend 64" "weekend 8" "cornell 32" "mesh 16" "exterior 16"; do set -- $sc; python tools/tune.py --scene $1 --spp $2 --reps 4 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1 $2', round(d['ms'],3), d['acc_sha'])"; done
