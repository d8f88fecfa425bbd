mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -k "dac" --timeout 120 --timeout-method=thread > gpurun_out/pytest_dac.log 2>&1; echo "pytest dac rc=$?"; tail -3 gpurun_out/pytest_dac.log
for impl in 0 3; do
timeout 120 python scripts/profile_decode.py --frames 4 --runs 3 --dac $impl --dac-frames 2580 > gpurun_out/dac_time_$impl.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/dac_time_$impl.log
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_dac_conv --csv --log-file gpurun_out/launches_dac_v3.csv python scripts/profile_decode.py --frames 4 --runs 1 --dac 3 --dac-frames 2580 > gpurun_out/ncu_dac_v3.log 2>&1; echo "ncu rc=$?"
