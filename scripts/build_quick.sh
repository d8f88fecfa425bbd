#!/bin/bash
# Developer rebuild: recompiles the translation units named on the command line (default: zb_api), relinks, stamps the source hash.
set -e
R=/root/repo/zonos_vibes_b200
for tu in ${@:-zb_api}; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -c $R/csrc/$tu.cu -o $R/csrc/$tu.o 2>&1 | grep -i "error" || true
done
nvcc -shared -o $R/libzonos_b200.so $R/csrc/zb_api.o $R/csrc/dac_api.o $R/csrc/spk_api.o -lcudart 2>/dev/null
python -m zonos_vibes_b200.build --stamp
