#!/bin/bash
# The checked build: every ZB_ASSERT (index / capacity assertions at the kernels' address computations, zb_common.cuh) compiled in.
# Output: zonos_vibes_b200/libzonos_b200_checked.so; run anything against it with ZB_LIB=$PWD/zonos_vibes_b200/libzonos_b200_checked.so.
set -e
R=/root/repo/zonos_vibes_b200
mkdir -p $R/csrc/_checked
pids=""
for tu in zb_api dac_api spk_api; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -DZB_BOUNDS -c $R/csrc/$tu.cu -o $R/csrc/_checked/$tu.o 2>&1 | grep -i "error" &
  pids="$pids $!"
done
wait
nvcc -shared -o $R/libzonos_b200_checked.so $R/csrc/_checked/zb_api.o $R/csrc/_checked/dac_api.o $R/csrc/_checked/spk_api.o -lcudart 2>/dev/null
ls -la $R/libzonos_b200_checked.so
