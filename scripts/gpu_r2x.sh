mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
