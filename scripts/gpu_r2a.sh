mkdir -p gpurun_out
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4 --prefix-len 1290"
timeout 300 $P3 > gpurun_out/plain3.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -f -o gpurun_out/prof_mega2 $P3 > gpurun_out/ncu3.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu3.log
