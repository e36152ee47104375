set -u
mkdir -p gpurun_out
run8() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/$1.json 2> gpurun_out/$1.err; echo "$1 rc=$?"; }
run8 sab_n8_direct
ZB_BENCH_DIRECT=0 run8 sab_n8_staged
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/sab_n1.json 2> gpurun_out/sab_n1.err; echo "n1 rc=$?"
python - <<'PY'
import json
for f in ('sab_n1','sab_n8_direct','sab_n8_staged'):
    try:
        d=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
        print(f, 'value %.0f ms %.3f | e2e %.0f ms %.3f %s registered %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], {k:round(v,3) for k,v in d['e2e']['parts_ms'].items()}, d['e2e'].get('output_registered')))
    except Exception as e: print(f, 'ERR', e)
PY
