#!/bin/bash
# Batch-64 decode step: warm-cache per-kernel durations (ncu --cache-control none; serialised, no PDL overlap), two steps at short
# context (the attention at ctx 1 316 is in r02_prof_attn_tma_raw.csv).  Plain run first.
mkdir -p gpurun_out
CMD="python scripts/profile_decode.py --frames 8 --runs 1 --batch 64 --prefix-len 24"
$CMD > gpurun_out/b64_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -k regex:"k_bt|k_attn|k_sample|k_update|k_embed" -s 900 -c 440 --csv --log-file gpurun_out/r02_b64_warm_kernels_final.csv $CMD > gpurun_out/ncu_b64.log 2>&1
echo "ncu rc=$?"; tail -1 gpurun_out/b64_plain.log; tail -2 gpurun_out/ncu_b64.log
