mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 900 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2r_bench_n8.json 2> gpurun_out/r2r_bench_n8.err
timeout 400 $TR tools/links_multi_gpu.py --events-per-shard 100000 --accounts 120000 --check --reps 2 > gpurun_out/r2r_links_n8_nccl_check.json 2> gpurun_out/r2r_links_n8_nccl.err
timeout 400 $TR tools/links_multi_gpu.py --events-per-shard 100000 --accounts 120000 --check --reps 2 --peer > gpurun_out/r2r_links_n8_peer_check.json 2> gpurun_out/r2r_links_n8_peer.err
tail -n 4 gpurun_out/r2r_bench_n8.err; cat gpurun_out/r2r_links_n8_nccl_check.json gpurun_out/r2r_links_n8_peer_check.json; cut -c 1-300 gpurun_out/r2r_bench_n8.json
