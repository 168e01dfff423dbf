mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > gpurun_out/r2i_gputest.log
timeout 900 python bench.py --steps 20 --warmup 5 --skip path_a,ltv --skip-cpu > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
timeout 600 python tools/profile_target.py c3 2 > gpurun_out/r2i_c3_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r2i_c3_launches.csv python tools/profile_target.py c3 2 > gpurun_out/r2i_c3_ncu.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2i_fused_launches.csv python tools/profile_target.py fused 2 > gpurun_out/r2i_fused_launches.log 2>&1
timeout 600 python tools/bench_fused_parts.py > gpurun_out/r2i_fused_parts.json 2>&1
tail -5 gpurun_out/r2i_gputest.log; tail -n 3 gpurun_out/r2i_bench.err; cat gpurun_out/r2i_c3_plain.log gpurun_out/r2i_fused_parts.json
