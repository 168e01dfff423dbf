mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.build(); g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2z_bench_reference.json 2> gpurun_out/r2z_bench_reference.err
timeout 1200 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err
tail -n 2 gpurun_out/r2z_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2z_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['two_slots_in_flight']['ms_per_step'], d['roofline']['traffic_stale'], d['clocks'])
for b in ('path_a','c3','live_loop','ltv'):
    x=d.get(b,{}); print(b, x.get('value') or x.get('tx_per_s_steady'), x.get('ms_per_step'), (x.get('roofline') or {}).get('frac'))
r=json.load(open('gpurun_out/r2z_bench_reference.json')); print('ref', r['value'])
PY
