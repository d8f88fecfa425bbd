mkdir -p gpurun_out
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=3 > gpurun_out/trace_att.log 2>&1; echo "trace rc=$?"
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 24 --frames 120 --runs 1 --opt trace=3 > gpurun_out/trace_att24.log 2>&1; echo "trace rc=$?"
