mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -30) > gpurun_out/r2f_gputest.log
(IGS_SORT_PAIRS=1 timeout 900 python -m pytest tests/test_velocity_gpu.py tests/test_append_events_gpu.py tests/test_identifiers_gpu.py -m gpu -q -x 2>&1 | tail -8) > gpurun_out/r2f_gputest_pairs.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err
timeout 600 python tools/bench_velocity.py > gpurun_out/r2f_velocity.json 2>&1
timeout 600 python tools/bench_fused_parts.py > gpurun_out/r2f_fused_parts.json 2>&1
tail -30 gpurun_out/r2f_gputest.log; tail -8 gpurun_out/r2f_gputest_pairs.log; tail -n 3 gpurun_out/r2f_bench.err; cat gpurun_out/r2f_fused_parts.json; tail -c 1200 gpurun_out/r2f_velocity.json
