#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_configs.py  -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
echo "=== prefill A/B (deferred max vs max first)"
for i in 1 2 3; do
timeout 200 python tools/probe_prefill.py 2>&1 | tail -2
CVLLM_LIB=compactor-vllm_b200/lib/variants/lib_maxfirst.so timeout 200 python tools/probe_prefill.py 2>&1 | tail -2
done
