# evidence for the epilogue-warp kernel: bench line, ncu launch list of the bench command, ncu --set full capture
mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_j.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4 --prefix-len 1290"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -f -o gpurun_out/prof_mega4 $P3 > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=2 > gpurun_out/trace_fine_epi.log 2>&1
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace_epi.log 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
