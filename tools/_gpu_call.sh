mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4) > gpurun_out/r3b_gputest.log
tail -2 gpurun_out/r3b_gputest.log
timeout 200 python tools/profile_target.py fused 2 2>&1 | tail -1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_score_ws|k_visit_keys|k_os_pass" -c 12 -o gpurun_out/r3b_fused -f python tools/profile_target.py fused 2 > gpurun_out/r3b_fused_ncu.log 2>&1
tail -n 1 gpurun_out/r3b_fused_ncu.log
