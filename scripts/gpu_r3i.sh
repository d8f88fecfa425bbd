# ncu evidence for the final kernel: launch list of the bench command + --set full capture (DRAM traffic)
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4 --prefix-len 1290"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -f -o gpurun_out/prof_mega6 $P3 > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
