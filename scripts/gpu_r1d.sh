# consumer-loop fix (LDS + independent HMMA accumulators) and ring-depth experiment
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
rm -f gpurun_out/probe.log
for o in "mega=1" "ring=3" "ring=2" "mega=0"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 --opt ring=2 > gpurun_out/trace_ring2.log 2>&1
tail -5 gpurun_out/pytest_all.log; cat gpurun_out/probe.log; grep prof gpurun_out/trace.log | tail -3
