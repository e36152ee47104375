cd /root/repo
python bench.py --steps 5 --warmup 3 2>/dev/null | tail -1 > gpurun_out/bench_pq.json
python -c "
import json; d=json.load(open('gpurun_out/bench_pq.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_bench_launches_pq.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_pq -s 1 -c 1 -f -o gpurun_out/k_render_pq python tools/tune.py --scene weekend --spp 64 --reps 1 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep
