mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for mma in 1 0; do
timeout 300 python scripts/profile_decode.py --batch 64 --prefix-len 1290 --frames 64 --runs 2 --opt attn_mma=$mma 2>&1 | tail -1
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 128 --runs 2 --opt attn_mma=$mma --opt mega=0 2>&1 | tail -1
done
timeout 300 python scripts/profile_decode.py --batch 64 --prefix-len 24 --frames 64 --runs 2 2>&1 | tail -1
