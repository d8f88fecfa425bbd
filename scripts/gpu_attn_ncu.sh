#!/bin/bash
# ncu --set full of the batched attention (k_attn_decode_tma) and the CTA-pair GEMM at batch 64, ctx ~1300; cold cache (default cache control:
# the KV of a layer is 351 MB, far beyond L2 anyway).  Plain run first.
mkdir -p gpurun_out
CMD="python scripts/profile_decode.py --frames 6 --runs 1 --batch 64 --prefix-len 1290"
$CMD > gpurun_out/attn_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:"k_attn_decode_tma" -s 33580 -c 2 -f -o gpurun_out/r02_attn_tma $CMD > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; tail -1 gpurun_out/attn_plain.log; tail -2 gpurun_out/ncu_attn.log
