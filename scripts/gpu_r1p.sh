mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "dac" --timeout 300 --timeout-method=thread > gpurun_out/pytest_dac.log 2>&1; echo "pytest dac rc=$?"; tail -3 gpurun_out/pytest_dac.log
timeout 300 python scripts/profile_decode.py --frames 4 --runs 3 --dac 0 --dac-frames 2580 2>&1 | tail -2
