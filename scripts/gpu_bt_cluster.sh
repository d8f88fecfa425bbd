#!/bin/bash
# fc1 (the one GEMM left on k_bt_gemm) with the activation tile multicast across clusters of 2 / 4 / 8 CTAs; batch 64, short context
mkdir -p gpurun_out
{
for o in "bt_cluster=1" "bt_cluster=2" "bt_cluster=4" "bt_cluster=8" "bt_cluster=1"; do
  echo "== $o"; timeout -s KILL 300 python scripts/profile_decode.py --batch 64 --frames 48 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
done
} | tee gpurun_out/bt_cluster8_ab.txt
