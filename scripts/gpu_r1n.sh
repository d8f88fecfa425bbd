mkdir -p gpurun_out
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -q -k "dac" --timeout 300 --timeout-method=thread > gpurun_out/pytest_dac.log 2>&1; echo "pytest dac rc=$?"; tail -15 gpurun_out/pytest_dac.log
timeout 300 python scripts/profile_decode.py --frames 4 --runs 1 --dac 0 --dac-frames 2580 2>&1 | tail -2
