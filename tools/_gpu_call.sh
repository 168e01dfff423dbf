mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4) > gpurun_out/r2t_gputest.log
tail -3 gpurun_out/r2t_gputest.log
timeout 120 python tools/bench_fused_parts.py 2>&1 | tail -1
timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv,c3,live_loop --skip-cpu > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err
tail -n 3 gpurun_out/r2t_bench.err
