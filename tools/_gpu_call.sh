mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR tools/links_multi_gpu.py --events-per-shard 150000 --accounts 60000 --check --reps 2 > gpurun_out/r2g_links_n2_nccl_check.json 2> gpurun_out/r2g_links_n2_nccl.err
timeout 600 $TR tools/links_multi_gpu.py --events-per-shard 150000 --accounts 60000 --check --reps 2 --peer > gpurun_out/r2g_links_n2_peer_check.json 2> gpurun_out/r2g_links_n2_peer.err
timeout 900 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2g_bench_n2.json 2> gpurun_out/r2g_bench_n2.err
(timeout 600 python -m pytest tests/test_append_events_gpu.py -m gpu -q -x 2>&1 | tail -5) > gpurun_out/r2g_gputest.log
cat gpurun_out/r2g_links_n2_nccl_check.json gpurun_out/r2g_links_n2_peer_check.json; tail -n 5 gpurun_out/r2g_links_n2_nccl.err gpurun_out/r2g_links_n2_peer.err gpurun_out/r2g_bench_n2.err; cut -c 1-400 gpurun_out/r2g_bench_n2.json; cat gpurun_out/r2g_gputest.log
