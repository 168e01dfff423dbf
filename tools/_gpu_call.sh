mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > gpurun_out/r2j_gputest.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv --skip-cpu > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
timeout 600 python tools/bench_velocity.py > gpurun_out/r2j_velocity.json 2>&1
timeout 600 python tools/bench_fused_parts.py > gpurun_out/r2j_fused_parts.json 2>&1
tail -5 gpurun_out/r2j_gputest.log; tail -n 3 gpurun_out/r2j_bench.err; cat gpurun_out/r2j_fused_parts.json; tail -c 900 gpurun_out/r2j_velocity.json
