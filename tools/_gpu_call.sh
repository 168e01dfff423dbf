mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --skip path_a,ltv,c3,live_loop --skip-cpu --skip-latency > gpurun_out/r3e_bench.json 2> gpurun_out/r3e_bench.err
tail -3 gpurun_out/r3e_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r3e_bench.json'))
print(d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['value'], d['e2e']['two_slots_in_flight'])
PY
