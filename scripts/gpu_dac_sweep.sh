#!/bin/bash
# DAC config 4 option sweep (utterance group size, tile options), then a sanity run of the persistent kernel.
mkdir -p gpurun_out
for o in "max_group=8" "max_group=4" "max_group=16" "max_group=2" "mt2_min_waves=1" "mt2_min_waves=3" "wide_cout=0" "wide_cout=1"; do
  timeout 120 python scripts/dac_bench.py --opt $o 2>&1 | tail -1
done | tee gpurun_out/dac_sweep.log
timeout 120 python scripts/profile_decode.py --frames 256 --runs 2 --prefix-len 1290 2>&1 | tail -1
