#!/bin/bash
# Batch-64 decode at ctx ~1300: warm-cache per-kernel durations (ncu --cache-control none; serialised, no PDL overlap).
mkdir -p gpurun_out
CMD="python scripts/profile_decode.py --frames 8 --runs 1 --batch 64 --prefix-len 1290"
$CMD > gpurun_out/b64_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none --cache-control none -k regex:"k_bt|k_attn|k_sample|k_update|k_embed" -s 900 -c 420 --csv --log-file gpurun_out/r02_b64_warm_kernels.csv $CMD > gpurun_out/ncu_b64.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/b64_plain.log; tail -3 gpurun_out/ncu_b64.log
