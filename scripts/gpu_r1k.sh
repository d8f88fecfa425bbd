mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 --timeout-method=thread > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
tail -40 gpurun_out/pytest_all.log
timeout 600 python scripts/profile_decode.py --frames 200 --runs 2 --batch 8 > gpurun_out/probe_b.log 2>&1
timeout 600 python scripts/profile_decode.py --frames 200 --runs 2 --batch 64 >> gpurun_out/probe_b.log 2>&1
timeout 600 python scripts/profile_decode.py --frames 1200 --runs 1 --batch 64 >> gpurun_out/probe_b.log 2>&1
cat gpurun_out/probe_b.log
