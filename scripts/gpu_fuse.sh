#!/bin/bash
# in_proj finish fused into the attention prologue: parity, then A/B at several batch sizes
mkdir -p gpurun_out
timeout -s KILL 500 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "batched or other_paths or audio_prefix or long_context or hybrid or engine or tma_staging" 2>&1 | tail -4
rc=${PIPESTATUS[0]}; echo "pytest rc=$rc"
{
for o in "attn_fuse=1" "attn_fuse=0"; do
  echo "== b64 short ctx $o"; timeout -s KILL 300 python scripts/profile_decode.py --batch 64 --frames 48 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
  echo "== b64 ctx 1300 $o"; timeout -s KILL 300 python scripts/profile_decode.py --batch 64 --frames 48 --runs 2 --prefix-len 1290 --opt $o 2>&1 | tail -1
  echo "== b16 short ctx $o"; timeout -s KILL 300 python scripts/profile_decode.py --batch 16 --frames 48 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
  echo "== b16 ctx 1300 $o"; timeout -s KILL 300 python scripts/profile_decode.py --batch 16 --frames 48 --runs 2 --prefix-len 1290 --opt $o 2>&1 | tail -1
  echo "== hybrid b16 $o"; timeout -s KILL 300 python scripts/profile_decode.py --hybrid --batch 16 --frames 96 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
done
} | tee gpurun_out/attn_fuse_ab.txt
