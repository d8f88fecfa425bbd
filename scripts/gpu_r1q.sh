mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "dac" --timeout 300 --timeout-method=thread > gpurun_out/pytest_dac.log 2>&1; echo "pytest dac rc=$?"; tail -3 gpurun_out/pytest_dac.log
timeout 300 python scripts/profile_decode.py --frames 4 --runs 3 --dac 0 --dac-frames 2580 > gpurun_out/dac_time.log 2>&1; rc=$?; tail -2 gpurun_out/dac_time.log
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_dac --csv --log-file gpurun_out/launches_dac_q.csv python scripts/profile_decode.py --frames 4 --runs 1 --dac 0 --dac-frames 2580 > gpurun_out/ncu_dac_q.log 2>&1; echo "ncu rc=$?"
fi
