# same-box A/B: lib_prev (committed) vs lib_new (packed MMA layout as a compile-time variant), then parity suite + a trace run on the new one
mkdir -p gpurun_out
L=zonos_vibes_b200/libzonos_b200.so
for rep in 1 2; do
cp zonos_vibes_b200/lib_prev.so.bin $L
echo "prev"; timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
cp zonos_vibes_b200/lib_new.so.bin $L
echo "new"; timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 2>&1 | tail -1
done
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 120 --runs 1 --opt trace=1 2>&1 | grep -c "skew"
