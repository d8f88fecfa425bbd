# 8-GPU check of the bench contract (one rank per GPU, weak scaling, no collective on the data path)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$?"; tail -1 gpurun_out/bench_n8.json | cut -c1-300; tail -3 gpurun_out/bench_n8.err
