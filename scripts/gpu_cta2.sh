#!/bin/bash
# k_bt_gemm2 (CTA pairs): parity, then which GEMMs gain (bit 0 = split-K GEMMs, bit 1 = fc1).  Batch 64, short context (GEMM-dominated) and ctx 1300.
mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "cta_pairs" 2>&1 | tail -15
rc=${PIPESTATUS[0]}; echo "pytest rc=$rc"
[ "$rc" != "0" ] && exit 0
for o in "bt_cta2=0" "bt_cta2=1" "bt_cta2=2" "bt_cta2=3"; do
  echo "== short ctx $o"
  timeout -s KILL 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch 64 --prefix-len 24 --opt $o 2>&1 | tail -1
done | tee -a gpurun_out/bt_cta2_ab.txt
for o in "bt_cta2=0" "bt_cta2=1" "bt_cta2=3"; do
  echo "== ctx 1300 $o"
  timeout -s KILL 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch 64 --prefix-len 1290 --opt $o 2>&1 | tail -1
done | tee -a gpurun_out/bt_cta2_ab.txt
