#!/bin/bash
# ncu --set full of the batched path's GEMMs (one decode layer's worth: in_proj, out_proj, fc1, fc2), warm cache, batch 64:
# k_bt_gemm (default) or, with argument 2, k_bt_gemm2 (CTA pairs).  Plain run first.
mkdir -p gpurun_out
if [ "$1" = "2" ]; then K='^k_bt_gemm2$'; OPT="--opt bt_cta2=3"; OUT=r02_bt_gemm2; else K='^k_bt_gemm$'; OPT=""; OUT=r02_bt_gemm; fi
CMD="python scripts/profile_decode.py --frames 6 --runs 1 --batch 64 --prefix-len 24 $OPT"
$CMD > gpurun_out/btg_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --cache-control none -k regex:"$K" -s 300 -c 4 -f -o gpurun_out/$OUT $CMD > gpurun_out/ncu_btg1.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/btg_plain.log; tail -3 gpurun_out/ncu_btg1.log
