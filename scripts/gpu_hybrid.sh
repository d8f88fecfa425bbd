#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "hybrid" 2>&1 | tail -6
timeout 300 python scripts/profile_decode.py --hybrid --batch 16 --frames 48 --runs 2 --prefix-len 24 2>&1 | tail -2
timeout 300 python scripts/profile_decode.py --hybrid --batch 16 --frames 48 --runs 2 --prefix-len 24 --opt bt_min_rows=999 2>&1 | tail -1
