mkdir -p gpurun_out
(timeout 200 python -m pytest tests/test_ltv_gpu.py -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r2w_ltv.log
tail -3 gpurun_out/r2w_ltv.log
timeout 120 python tools/profile_target.py ltv 6 2>&1 | tail -1
timeout 600 python bench.py --steps 5 --warmup 3 --skip path_a,live_loop --skip-cpu > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err
tail -n 2 gpurun_out/r2w_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2w_bench.json'))
print('fused', d['ms_per_step'], 'c3', {k:v for k,v in d['c3'].items() if k in ('build','parts_ms','ms_per_step','value')} )
print('ltv', d['ltv'].get('ms_per_step'), d['ltv'].get('roofline',{}).get('frac'))
PY
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_part -c 12 --csv --log-file gpurun_out/r2w_part_launches.csv python tools/bench_partition.py > gpurun_out/r2w_part_ncu.log 2>&1
tail -n 8 gpurun_out/r2w_part_launches.csv
