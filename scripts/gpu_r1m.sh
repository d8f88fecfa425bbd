mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
P="python scripts/profile_decode.py --frames 4 --runs 1 --dac 0 --dac-frames 2580"
timeout 300 $P > gpurun_out/plain_dac.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:dac --csv --log-file gpurun_out/launches_dac.csv $P > gpurun_out/ncu_dac.log 2>&1
tail -3 gpurun_out/plain_dac.log
