mkdir -p gpurun_out
timeout 120 python scripts/profile_decode.py --frames 4 --runs 1 --dac 0 --dac-frames 2580 > gpurun_out/dac_time_v2.log 2>&1; rc=$?; tail -1 gpurun_out/dac_time_v2.log
if [ $rc -eq 0 ]; then
  timeout 800 ncu --set full --import-source on --clock-control none -k regex:k_dac_conv_tc2 --launch-skip 16 --launch-count 2 -f -o gpurun_out/r01_dac_v2_l3 python scripts/profile_decode.py --frames 4 --runs 1 --dac 0 --dac-frames 2580 > gpurun_out/ncu_dac_full.log 2>&1; echo "ncu rc=$?"
fi
