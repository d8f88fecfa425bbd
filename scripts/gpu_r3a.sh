# final round-1 run: parity suite (with the epoch-tagged-vs-barrier test), smoke, bench line, att_chunk sweep
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err; echo "bench rc=$?"; python3 -c "
import json; d=json.load(open('gpurun_out/bench_f.json')); print(d['value'], d['us_per_decode_step'], d['roofline']['frac'], d['roofline']['traffic'], d['e2e']['value'], d['gpu_launches'], d['clocks'])"
for ch in 64 96 128; do
for pl in 600 1290 2300; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len $pl --frames 256 --runs 2 --opt att_chunk=$ch 2>&1 | tail -1
done
done
