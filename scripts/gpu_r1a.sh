# Round-1 re-baseline: GPU parity tests, bench line, ncu launch list of a short full-size generation.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 --timeout-method=thread --durations=15 > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/bench.err
rm -f gpurun_out/probe.log
for o in "gemv=1" "gemv=0" "pdl=0" "graph=0"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 64 --runs 2 --dac 0 --dac-frames 2580 >> gpurun_out/probe.log 2>&1
timeout 600 python scripts/profile_decode.py --frames 16 --runs 2 --opt graph=0 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -s 6600 -c 420 --csv --log-file gpurun_out/launches.csv python scripts/profile_decode.py --frames 16 --runs 2 --opt graph=0 > gpurun_out/ncu1.log 2>&1
tail -4 gpurun_out/pytest_all.log; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err; cat gpurun_out/probe.log
