mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_cpp_host.py -m gpu -q -x 2>&1 | tail -2)
g++ -std=c++17 -O2 -Wall -pthread -I include tools/sharded_engine_run.cc -L igaming-platform_b200 -ligs -Wl,-rpath,$PWD/igaming-platform_b200 -o gpurun_out/sharded_engine_run || exit 1
nproc
timeout 300 gpurun_out/sharded_engine_run 1 4000000 | tee gpurun_out/r2z_sharded_engine_n1.json | cut -c1-250
timeout 300 gpurun_out/sharded_engine_run 2 4000000 | tee gpurun_out/r2z_sharded_engine_n2.json | cut -c1-250
timeout 300 gpurun_out/sharded_engine_run 2 16000000 | tee gpurun_out/r2z_sharded_engine_n2_16m.json | cut -c1-250
