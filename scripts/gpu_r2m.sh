mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for ch in 256 192 128 64; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 --opt att_chunk=$ch 2>&1 | tail -1
done
for ch in 256 192 128; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 600 --frames 256 --runs 2 --opt att_chunk=$ch 2>&1 | tail -1
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 2300 --frames 256 --runs 2 --opt att_chunk=$ch 2>&1 | tail -1
done
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=3 > gpurun_out/trace_att.log 2>&1; echo "trace rc=$?"
