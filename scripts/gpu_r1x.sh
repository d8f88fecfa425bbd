mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace_1290.log 2>&1; echo "trace rc=$?"
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 24 --frames 120 --runs 1 --opt trace=1 > gpurun_out/trace_24.log 2>&1
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
timeout 300 python scripts/profile_decode.py --batch 1 --prefix-len 24 --frames 256 --runs 2 2>&1 | tail -1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ring_bench scripts/ring_bench.cu && timeout 120 /tmp/ring_bench > gpurun_out/ring_bench.txt 2>&1; tail -8 gpurun_out/ring_bench.txt
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err; echo "bench rc=$?"; cat gpurun_out/bench_x.json
