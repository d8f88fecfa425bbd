#!/bin/bash
# A/B of the batched attention's K/V staging (ld.global fragments vs cp.async.bulk ring, slot released after / before the math; ring depth) at batch 64, ctx ~1300-1350.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "tma_attn or tma_staging or equals_independent" 2>&1 | tail -3
for o in "attn_tma=0" "attn_tma=1" "attn_tma=1 --opt attn_stages=2" "attn_tma=1 --opt attn_stages=4" "attn_tma=2 --opt attn_stages=2"; do
  echo "== $o"
  timeout 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch 64 --prefix-len 1290 --opt $o 2>&1 | tail -1
done | tee gpurun_out/attn_tma_ab2.txt
echo "== batch 16 / 32"
for b in 16 32; do for o in "attn_tma=0" "attn_tma=1" "attn_tma=1 --opt attn_stages=2"; do
  timeout 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch $b --prefix-len 1290 --opt $o 2>&1 | tail -1
done; done | tee -a gpurun_out/attn_tma_ab2.txt
