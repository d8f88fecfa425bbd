#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/profile_decode.py --hybrid --batch 16 --frames 6 --runs 1 --prefix-len 24"
$CMD > gpurun_out/hyb_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -k regex:"k_bt_|k_m2_ssm|k_m2_conv\(|k_m2_gated|k_attn_decode" -s 300 -c 420 --csv --log-file gpurun_out/r02_hybrid_b16_kernels.csv $CMD > gpurun_out/ncu_hyb.log 2>&1
echo "ncu rc=$?"; tail -1 gpurun_out/hyb_plain.log
