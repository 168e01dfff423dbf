mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --skip path_a,live_loop,c3 --skip-cpu > gpurun_out/r2z_bench_n2.json 2> gpurun_out/r2z_bench_n2.err
echo rc=$?
tail -n 3 gpurun_out/r2z_bench_n2.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2z_bench_n2.json'))
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'])
c=d.get('c5',{}); print(c.get('value'), c.get('ms_per_step'), c.get('parts_ms'), c.get('partition_roofline',{}).get('frac'))
print(d.get('ltv',{}).get('ms_per_step'))
PY
