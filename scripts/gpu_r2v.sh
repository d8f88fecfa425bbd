mkdir -p gpurun_out
timeout 100 python scripts/profile_decode.py --batch 1 --prefix-len 24 --frames 32 --runs 1 2>&1 | tail -1
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for rep in 1 2; do
for l in 1 0; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 --opt ll=$l 2>&1 | tail -1
done
done
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 24 --frames 256 --runs 2 2>&1 | tail -1
timeout 200 python scripts/profile_decode.py --batch 2 --prefix-len 1290 --frames 128 --runs 2 2>&1 | tail -1
timeout 200 python scripts/profile_decode.py --batch 4 --prefix-len 1290 --frames 128 --runs 2 2>&1 | tail -1
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace_ll.log 2>&1; echo "trace rc=$?"
