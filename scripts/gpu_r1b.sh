# megakernel bring-up: parity tests + short timing probes (old path vs persistent kernel) + phase trace
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 --timeout-method=thread > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
rm -f gpurun_out/probe.log
for o in "pf=10" "pf=0" "pf=4" "pf=20" "mega=0"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 2580 --runs 1 >> gpurun_out/probe.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace.log 2>&1
tail -5 gpurun_out/pytest_all.log; cat gpurun_out/probe.log
