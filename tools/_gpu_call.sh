mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 --skip path_a,live_loop,c3 --skip-cpu > gpurun_out/r3d_bench_n8.json 2> gpurun_out/r3d_bench_n8.err
echo rc=$?
python - <<'PY'
import json
lines=[l for l in open('gpurun_out/r3d_bench_n8.json') if l.startswith('{')]
d=json.loads(lines[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
c=d.get('c5',{}); print(c.get('value'), c.get('ms_per_step'), c.get('parts_ms'), c.get('partition_roofline',{}).get('frac'))
print(d.get('ltv',{}).get('ms_per_step'), d.get('ltv',{}).get('value'))
PY
