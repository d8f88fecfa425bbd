mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 240 --timeout-method=thread > gpurun_out/pytest_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_all.log
for o in "pdl=1" "pdl=0" "graph=0" "l2_prefetch=16" "l2_prefetch=48"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 64 --runs 1 --dac 0 --dac-frames 2580 >> gpurun_out/probe.log 2>&1
timeout 300 python scripts/profile_decode.py --frames 8 --runs 1 --dac 1 --dac-frames 2580 >> gpurun_out/probe.log 2>&1
timeout 600 python scripts/profile_decode.py --frames 16 --runs 2 --opt graph=0 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -s 6600 -c 420 --csv --log-file gpurun_out/launches.csv python scripts/profile_decode.py --frames 16 --runs 2 --opt graph=0 > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:_Z6k_gemvILi2ELi2E -s 700 -c 2 -o gpurun_out/prof_fc1 python scripts/profile_decode.py --frames 16 --runs 2 --opt graph=0 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/pytest_all.log; cat gpurun_out/probe.log
