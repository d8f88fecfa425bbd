# round-1 evidence run: parity suite, stage-level fc1 trace, bench line, ncu launch list of the bench command,
# ncu --set full capture of the persistent kernel (DRAM traffic), DAC config 4, batch-64 probe
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=2 > gpurun_out/trace_fine8.log 2>&1; echo "trace rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_e.json 2> gpurun_out/bench_e.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_e.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4 --prefix-len 1290"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -f -o gpurun_out/prof_mega3 $P3 > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
timeout 600 python scripts/dac_ab.py --frames 2580 --batch 32 --reps 3 base > gpurun_out/dac_b32.txt 2>&1; cat gpurun_out/dac_b32.txt | tail -2
timeout 600 python scripts/dac_ab.py --frames 2580 --batch 1 --reps 5 base > gpurun_out/dac_b1.txt 2>&1; cat gpurun_out/dac_b1.txt | tail -2
timeout 600 python scripts/profile_decode.py --batch 64 --prefix-len 1290 --frames 48 --runs 2 2>&1 | tail -1
timeout 600 python scripts/profile_decode.py --batch 64 --prefix-len 24 --frames 48 --runs 2 2>&1 | tail -1
timeout 600 python scripts/profile_decode.py --batch 16 --prefix-len 1290 --frames 48 --runs 2 2>&1 | tail -1
