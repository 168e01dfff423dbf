mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40) > gpurun_out/r2c_gputest.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
tail -40 gpurun_out/r2c_gputest.log; tail -n 3 gpurun_out/r2c_bench.err; cut -c 1-300 gpurun_out/r2c_bench.json
