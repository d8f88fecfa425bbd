#!/bin/bash
# Batched tcgen05 path: activation tile multicast across a cluster of 1 / 2 / 4 CTAs; parity subset first (under a timeout).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "batched or other_paths or long_context or audio_prefix or hybrid or plugin" 2>&1 | tail -4
for cs in 1 2 4; do
  timeout 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch 64 --prefix-len 24 --opt bt_cluster=$cs 2>&1 | tail -1
done | tee gpurun_out/bt_cluster_ab.log
