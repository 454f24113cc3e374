#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scores.py tests/test_gpu_configs.py tests/test_gpu_engine.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
echo "=== scoring stages"
timeout 300 python tools/prof_scoring.py 2>&1 | tail -5
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/scoring_launches.csv python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
grep -E "snap_" gpurun_out/scoring_launches.csv | awk -F'","' '{print substr($5,1,60), $NF}' | tail -6
