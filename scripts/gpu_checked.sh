#!/bin/bash
# The whole GPU suite against the checked build (ZB_ASSERT compiled in): the substitute for compute-sanitizer memcheck on this pool.
mkdir -p gpurun_out
ZB_LIB=$PWD/zonos_vibes_b200/libzonos_b200_checked.so timeout -s KILL 900 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_checked.log 2>&1; echo "checked suite rc=$?"
tail -3 gpurun_out/pytest_checked.log; grep -c "ZB_ASSERT failed" gpurun_out/pytest_checked.log
ZB_LIB=$PWD/zonos_vibes_b200/libzonos_b200_checked.so timeout -s KILL 300 python scripts/profile_decode.py --frames 32 --runs 1 --batch 64 --prefix-len 300 2>&1 | tail -1
ZB_LIB=$PWD/zonos_vibes_b200/libzonos_b200_checked.so timeout -s KILL 300 python scripts/profile_decode.py --frames 64 --runs 1 --batch 1 --prefix-len 2500 2>&1 | tail -1
