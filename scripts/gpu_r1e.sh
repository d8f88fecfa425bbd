# no local memory in the consumer loops, adaptive KV splits, page table in registers; ring-depth sweep
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
rm -f gpurun_out/probe.log
for o in "ring=5" "ring=4" "ring=3" "ring=2"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 2580 --runs 1 --opt ring=3 >> gpurun_out/probe.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 2580 --runs 1 --opt ring=5 >> gpurun_out/probe.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 --opt ring=3 > gpurun_out/trace.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 --opt ring=5 > gpurun_out/trace_ring5.log 2>&1
tail -5 gpurun_out/pytest_all.log; cat gpurun_out/probe.log; grep prof gpurun_out/trace.log | tail -3
