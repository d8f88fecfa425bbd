# evidence for the x_ll kernel: parity suite, smoke, bench line, ncu launch list of the bench command, ncu --set full capture, traces
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_u.json 2> gpurun_out/bench_u.err; echo "bench rc=$?"; python3 -c "
import json; d=json.load(open('gpurun_out/bench_u.json')); print(d['value'], d['us_per_decode_step'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches'], d['clocks'])"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
P3="python scripts/profile_decode.py --frames 16 --runs 1 --steps-per-poll 4 --prefix-len 1290"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 3 -c 1 -f -o gpurun_out/prof_mega5 $P3 > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace_ll.log 2>&1
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=3 > gpurun_out/trace_att.log 2>&1
