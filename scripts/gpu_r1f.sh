# ring micro-benchmark sweep + persistent kernel: polling style A/B, LN single round trip, attention sub-stamps
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ring_bench scripts/ring_bench.cu && timeout 120 /tmp/ring_bench > gpurun_out/ring_bench.log 2>&1
rm -f gpurun_out/probe.log
for o in "ring=5" "ring=3" "poll_all=1"; do
  timeout 300 python scripts/profile_decode.py --frames 300 --runs 2 --opt $o >> gpurun_out/probe.log 2>&1
done
timeout 300 python scripts/profile_decode.py --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace.log 2>&1
cat gpurun_out/ring_bench.log; cat gpurun_out/probe.log; grep prof gpurun_out/trace.log | tail -3
