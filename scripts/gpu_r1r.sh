mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -k "dac" --timeout 120 --timeout-method=thread > gpurun_out/pytest_dac.log 2>&1; echo "pytest dac rc=$?"; tail -15 gpurun_out/pytest_dac.log
timeout 120 python scripts/profile_decode.py --frames 4 --runs 3 --dac 0 --dac-frames 2580 > gpurun_out/dac_time_v2.log 2>&1; echo "v2 rc=$?"; tail -2 gpurun_out/dac_time_v2.log
timeout 120 python scripts/profile_decode.py --frames 4 --runs 3 --dac 2 --dac-frames 2580 > gpurun_out/dac_time_v1.log 2>&1; echo "v1 rc=$?"; tail -1 gpurun_out/dac_time_v1.log
