#!/bin/bash
# A/B of the persistent kernel's dependency modes (ll = 0 grid barriers .. 3 every activation epoch-tagged; chain = MMA warps
# do not wait for their epilogue warp between phases), under a timeout (a mis-tagged word would spin forever), and a phase
# trace of the default mode.
mkdir -p gpurun_out
for o in "ll=0" "ll=1" "ll=2" "ll=3" "chain=1"; do
  timeout 120 python scripts/profile_decode.py --frames 256 --runs 2 --prefix-len 1290 --opt $o 2>&1 | tail -1
done | tee gpurun_out/ll_ab.log
timeout 120 python scripts/profile_decode.py --frames 64 --runs 1 --prefix-len 1290 --opt trace=1 > gpurun_out/phase_trace_ll3.log 2>&1
