#!/bin/bash
# small-batch checks of the two new batched-path defaults: hybrid batch 16 (32 rows) and transformer batch 8 / 16
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "batched or other_paths or audio_prefix or long_context or hybrid or engine" 2>&1 | tail -3
{
for o in "bt_cta2=1" "bt_cta2=0"; do
  echo "== hybrid b16 $o"
  timeout -s KILL 300 python scripts/profile_decode.py --hybrid --batch 16 --frames 96 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
done
for b in 8 16 32; do for o in "bt_cta2=1" "bt_cta2=0"; do
  echo "== transformer b$b short ctx $o"
  timeout -s KILL 300 python scripts/profile_decode.py --batch $b --frames 48 --runs 2 --prefix-len 24 --opt $o 2>&1 | tail -1
done; done
} | tee gpurun_out/small_batch_ab.txt
