mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_full_size_gpu.py tests/test_score_batch_gpu.py tests/test_append_events_gpu.py -m gpu -q -x 2>&1 | tail -4) > gpurun_out/r2l_gputest.log
tail -3 gpurun_out/r2l_gputest.log
for ws in 1 2 3 4 5; do for so in 3 0; do echo "WS=$ws SORT=$so"; IGS_SCORE_WS=$ws IGS_VISIT_SORT_PASSES=$so timeout 300 python tools/bench_fused_parts.py 2>&1 | tail -1; done; done > gpurun_out/r2l_ab.log 2>&1
cat gpurun_out/r2l_ab.log
(IGS_SCORE_WS=2 timeout 600 python -m pytest tests/test_full_size_gpu.py -m gpu -q -x -k c5 2>&1 | tail -3)
