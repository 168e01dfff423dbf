mkdir -p gpurun_out
for m in 1 2; do IGS_PART_MODE=$m timeout 300 python tools/bench_partition.py 2>&1 | tail -1 | cut -c1-40,230-; done
(timeout 300 python -m pytest tests -m gpu -q -x -k "partition or no_alloc or shard" 2>&1 | tail -2)
