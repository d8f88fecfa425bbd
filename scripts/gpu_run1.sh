mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 240 --timeout-method=thread -k "not dac" 2>&1 | tail -60 > gpurun_out/pytest_core.log
echo "core rc=$?" >> gpurun_out/pytest_core.log
timeout 300 python -m pytest tests -m gpu -q --timeout 120 --timeout-method=thread -k "dac_simt" 2>&1 | tail -30 > gpurun_out/pytest_dac_simt.log
timeout 300 python -m pytest tests -m gpu -q --timeout 90 --timeout-method=thread -k "dac_tcgen05 or dac_intermediates or dac_window" 2>&1 | tail -40 > gpurun_out/pytest_dac_tc.log
timeout 600 python bench.py --steps 1 --warmup 1 --dac-impl 1 --no-cpu-baseline > gpurun_out/bench_simt.log 2>&1
echo "bench rc=$?" >> gpurun_out/bench_simt.log
tail -5 gpurun_out/pytest_core.log gpurun_out/pytest_dac_simt.log gpurun_out/pytest_dac_tc.log gpurun_out/bench_simt.log
