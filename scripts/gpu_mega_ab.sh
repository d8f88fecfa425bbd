#!/bin/bash
# Persistent-kernel probe: us per decode step at ctx ~1290..1550 (3 runs), optional step-tail trace, then the gpu tests.
mkdir -p gpurun_out
python scripts/profile_decode.py --frames 256 --runs 3 --prefix-len 1290 "$@" 2>&1 | tee gpurun_out/mega_ab.log
