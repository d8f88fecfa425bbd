mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for sg in 0 256 384 512; do
for pl in 24 200 400; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len $pl --frames 128 --runs 2 --opt att_single=$sg 2>&1 | tail -1
done
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_o.json 2> gpurun_out/bench_o.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_o.json; python3 -c "
import json; d=json.load(open('gpurun_out/bench_o.json')); print(d['value'], d['us_per_decode_step'], d['roofline']['frac'])"
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --opt att_chunk=256 --opt att_single=256 > gpurun_out/bench_o2.json 2> gpurun_out/bench_o2.err; python3 -c "
import json; d=json.load(open('gpurun_out/bench_o2.json')); print(d['value'], d['us_per_decode_step'], d['roofline']['frac'])"
