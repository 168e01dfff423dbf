mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_score_batch_gpu.py tests/test_two_engines_gpu.py -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r3a_test.log
tail -5 gpurun_out/r3a_test.log
for d in 2 1; do
IGS_SCORE_DIRECT=$d timeout 600 python bench.py --steps 10 --warmup 3 --skip path_a,ltv,c3,live_loop --skip-cpu --skip-latency > gpurun_out/r3a_bench_d$d.json 2> gpurun_out/r3a_bench_d$d.err
python - <<PY
import json
d=json.load(open('gpurun_out/r3a_bench_d$d.json'))
print('direct=$d', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['value'])
PY
done
