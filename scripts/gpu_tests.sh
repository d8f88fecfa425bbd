#!/bin/bash
# GPU test pass: the no-alloc test alone in a fresh process, then the whole gpu suite (prints kept).
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu -s > gpurun_out/pytest_noalloc.log 2>&1; echo "noalloc rc=$?"
python -m pytest tests -q -m gpu -s --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?"
tail -40 gpurun_out/pytest_gpu.log
grep -E "^\[" gpurun_out/pytest_gpu.log | head -60
