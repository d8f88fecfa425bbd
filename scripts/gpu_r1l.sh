mkdir -p gpurun_out
timeout 600 python scripts/profile_decode.py --frames 200 --runs 2 --hybrid > gpurun_out/probe_h.log 2>&1
timeout 600 python scripts/profile_decode.py --frames 100 --runs 2 --hybrid --batch 16 >> gpurun_out/probe_h.log 2>&1
cat gpurun_out/probe_h.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_r01_n2.json 2> gpurun_out/bench_r01_n2.err
echo "n2 rc=$?"; cat gpurun_out/bench_r01_n2.json | cut -c1-600; tail -3 gpurun_out/bench_r01_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 3 --warmup 3 --impl reference > gpurun_out/bench_r01_n2_ref.json 2>> gpurun_out/bench_r01_n2.err
echo "n2 ref rc=$?"; cat gpurun_out/bench_r01_n2_ref.json | cut -c1-400
