#!/bin/bash
# Round evidence, call B: ncu --set full of ONE persistent-kernel launch (4 decode steps at ctx ~1300), plain run first.
mkdir -p gpurun_out
CMD="python scripts/profile_decode.py --frames 12 --runs 1 --prefix-len 1290 --steps-per-poll 4"
$CMD > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_decode_mega -s 1 -c 1 -f -o gpurun_out/r02_prof_mega $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/prof_plain.log; tail -5 gpurun_out/ncu_full.log
