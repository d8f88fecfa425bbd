mkdir -p gpurun_out
for ps in 100 0 20 50 200; do
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 --opt psleep=$ps 2>&1 | tail -1
done
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 --opt poll_all=1 2>&1 | tail -1
timeout 200 python scripts/profile_decode.py --batch 1 --prefix-len 1290 --frames 256 --runs 2 --opt psleep=0 --opt poll_all=1 2>&1 | tail -1
