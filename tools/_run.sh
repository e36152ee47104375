cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/tune.py --scene mesh --spp 64 --reps 3 2>/dev/null | tail -1 | cut -c1-330
timeout 900 python tools/run_c5.py 2>&1 | tail -3 | cut -c1-600
