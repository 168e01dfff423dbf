mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv,c3,live_loop --skip-cpu --skip-latency > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err
tail -n 3 gpurun_out/r2q_bench.err
IGS_SCORE_DIRECT=0 timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv,c3,live_loop --skip-cpu --skip-latency > gpurun_out/r2q_bench_staged.json 2> gpurun_out/r2q_bench2.err
(timeout 600 python -m pytest tests/test_score_batch_gpu.py tests/test_full_size_gpu.py tests/test_identifiers_gpu.py tests/test_two_engines_gpu.py -m gpu -q -x 2>&1 | tail -3)
