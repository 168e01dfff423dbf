mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4) > gpurun_out/r2n_gputest.log
tail -3 gpurun_out/r2n_gputest.log
for ws in 1 0; do for so in 3 0; do echo "WS=$ws SORT=$so"; IGS_SCORE_WS=$ws IGS_VISIT_SORT_PASSES=$so timeout 90 python tools/bench_fused_parts.py 2>&1 | tail -1; done; done > gpurun_out/r2n_ab.log 2>&1
cat gpurun_out/r2n_ab.log
timeout 300 python tools/bench_velocity.py 2>&1 | tail -c 700
