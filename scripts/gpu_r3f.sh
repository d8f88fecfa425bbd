# final check of the tree as committed: parity suite, smoke, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench rc=$?"; python3 -c "
import json; d=json.load(open('gpurun_out/bench_g.json')); print(d['value'], d['us_per_decode_step'], d['roofline']['frac'], d['roofline']['traffic'], d['e2e']['value'], d['gpu_launches'], d['clocks'])"
