mkdir -p gpurun_out
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=3 --opt att_chunk=256 > gpurun_out/trace_att.log 2>&1; echo "trace rc=$?"
