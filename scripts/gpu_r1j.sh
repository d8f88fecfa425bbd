# round-1 evidence: bench line, ncu launch lists (per-op batch 1, batched tcgen05 batch 64), ncu --set full captures
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --opt mega=0 --no-cpu-baseline > gpurun_out/bench_r01_perop.json 2>> gpurun_out/bench_r01.err; echo "bench perop rc=$?"
P1="python scripts/profile_decode.py --frames 8 --runs 1 --opt mega=0"
timeout 300 $P1 > gpurun_out/plain1.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 300 --csv --log-file gpurun_out/launches_perop_b1.csv $P1 > gpurun_out/ncu1.log 2>&1
P2="python scripts/profile_decode.py --frames 4 --runs 1 --batch 64"
timeout 300 $P2 > gpurun_out/plain2.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 7000 -c 450 --csv --log-file gpurun_out/launches_bt_b64.csv $P2 > gpurun_out/ncu2.log 2>&1
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4"
timeout 300 $P3 > gpurun_out/plain3.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -o gpurun_out/prof_mega $P3 > gpurun_out/ncu3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemv_mma -s 3100 -c 5 -o gpurun_out/prof_gemv $P1 > gpurun_out/ncu4.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_bt_gemm|k_attn_decode" -s 3000 -c 6 -o gpurun_out/prof_bt $P2 > gpurun_out/ncu5.log 2>&1
cat gpurun_out/bench_r01.json; tail -3 gpurun_out/bench_r01.err; tail -2 gpurun_out/ncu*.log
