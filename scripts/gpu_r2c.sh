mkdir -p gpurun_out
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=2 > gpurun_out/trace_fine.log 2>&1; echo "trace rc=$?"
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
