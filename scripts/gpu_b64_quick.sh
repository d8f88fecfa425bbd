#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "batched or other_paths or audio_prefix or long_context" 2>&1 | tail -3
timeout 200 python scripts/profile_decode.py --frames 48 --runs 2 --batch 64 --prefix-len 24 2>&1 | tail -1
