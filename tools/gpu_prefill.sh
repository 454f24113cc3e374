#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -q -x > gpurun_out/pytest_prefill.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_prefill.log
timeout 300 python tools/perf_probe.py > gpurun_out/perf_probe.log 2>&1; echo "probe rc=$?"; cat gpurun_out/perf_probe.log | tail -12
CVLLM_PREFILL_V1=1 timeout 300 python tools/perf_probe.py 2>&1 | grep prefill
