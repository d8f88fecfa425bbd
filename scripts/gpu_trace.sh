mkdir -p gpurun_out
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace.log 2>&1
echo "rc=$?" >> gpurun_out/trace.log
