#!/bin/bash
# Round evidence, call A: the default bench line, then the ncu launch list of a short bench command (plain run first).
mkdir -p gpurun_out
python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r02_bench.json
CMD="python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline"
$CMD > gpurun_out/r02_bench_short.json 2> gpurun_out/r02_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 1200 --csv --log-file gpurun_out/r02_launches_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?"
