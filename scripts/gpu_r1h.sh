mkdir -p gpurun_out
rm -f gpurun_out/probe.log
for o in "psleep=0" "psleep=100" "psleep=400" "psleep=1000"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace.log 2>&1
cat gpurun_out/probe.log; grep prof gpurun_out/trace.log | tail -3
