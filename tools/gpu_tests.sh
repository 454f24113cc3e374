#!/usr/bin/env bash
# Run every GPU test module in its own process (a CUDA fault in one must not hide the others).
# Usage (on the GPU box, from the repo root): bash tools/gpu_tests.sh [module ...]
set -u
mkdir -p gpurun_out
mods=("$@")
if [ ${#mods[@]} -eq 0 ]; then
  mods=(tests/test_gpu_selftest.py tests/test_gpu_store.py tests/test_gpu_select.py tests/test_gpu_decode.py tests/test_gpu_prefill.py tests/test_gpu_scores.py tests/test_gpu_attention_module.py)
fi
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
rc_all=0
for m in "${mods[@]}"; do
  [ -f "$m" ] || continue
  name=$(basename "$m" .py)
  echo "=== $m"
  timeout 900 python -m pytest "$m" -q -m gpu -x --no-header -p no:cacheprovider > "gpurun_out/$name.log" 2>&1
  rc=$?
  tail -n 3 "gpurun_out/$name.log"
  [ $rc -ne 0 ] && rc_all=1 && grep -E "^(FAILED|ERROR)|Error|error:|assert" "gpurun_out/$name.log" | head -20
done
exit $rc_all
