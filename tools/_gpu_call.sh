mkdir -p gpurun_out
timeout 200 python tools/profile_target.py c3 2 2>&1 | tail -1 || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 80 --csv --log-file gpurun_out/r2z_c3_launches.csv python tools/profile_target.py c3 2 > gpurun_out/r2z_c3_ncu.log 2>&1
tail -n 1 gpurun_out/r2z_c3_ncu.log
