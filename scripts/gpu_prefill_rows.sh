#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "several_row_tiles or batched or engine or audio_prefix or independent or plugin" 2>&1 | tail -4
# time to first frame, batch 64, 100-token prefix: generate 2 frames
timeout -s KILL 300 python scripts/profile_decode.py --batch 64 --frames 2 --runs 2 --prefix-len 100 2>&1 | tail -2
timeout -s KILL 300 python scripts/profile_decode.py --batch 16 --frames 2 --runs 2 --prefix-len 100 2>&1 | tail -1
