mkdir -p gpurun_out
(IGS_MLP_PAIRS=1 timeout 200 python -m pytest tests/test_ltv_gpu.py -m gpu -q -x 2>&1 | tail -25) > gpurun_out/r2v_ltv_pairs.log
tail -25 gpurun_out/r2v_ltv_pairs.log
IGS_MLP_PAIRS=0 timeout 120 python tools/profile_target.py ltv 6 2>&1 | tail -1
IGS_MLP_PAIRS=1 timeout 120 python tools/profile_target.py ltv 6 2>&1 | tail -1
