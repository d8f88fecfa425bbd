#!/bin/bash
# compute-sanitizer pass (ONE tool per gpurun call): tiny-config tests of the persistent kernel (ll=1 and ll=0), the
# per-op and batched tcgen05 paths, the DAC tcgen05 kernels and the chunked-scan prefill.   usage: gpu_sanitizer.sh <tool>
TOOL=${1:-memcheck}
mkdir -p gpurun_out
K='teacher_forced_logits and tiny or greedy_stream_identical and tiny or epoch_tagged and tiny or dac_tcgen05_matches_oracle and b1 or chunked_scan_prefill and 25 or audio_prefix and 5'
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 3 --print-limit 30 python -m pytest tests -q -m gpu -x -k "$K" > gpurun_out/sanitizer_$TOOL.log 2>&1
echo "sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|passed|failed|Error|RACECHECK SUMMARY" gpurun_out/sanitizer_$TOOL.log | tail -12
