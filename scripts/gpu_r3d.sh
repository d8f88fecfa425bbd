mkdir -p gpurun_out
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=4 > gpurun_out/trace_tail.log 2>&1; echo "trace rc=$?"
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
