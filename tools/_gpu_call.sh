mkdir -p gpurun_out
timeout 300 python tools/bench_partition.py 2>&1 | tail -1
(timeout 300 python -m pytest tests -m gpu -q -x -k "partition or no_alloc" 2>&1 | tail -3)
timeout 120 python tools/profile_target.py ltv 2 2>&1 | tail -1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mlp_gemm -c 6 -o gpurun_out/r2v_mlp -f python tools/profile_target.py ltv 2 > gpurun_out/r2v_mlp_ncu.log 2>&1
tail -2 gpurun_out/r2v_mlp_ncu.log
