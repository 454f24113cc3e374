#!/bin/bash
# leverage chain: unit tests, self-distance vs the reference, chain timing, per-kernel ncu durations
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scores.py -x -q -k "leverage or eigensolver or compactor" > gpurun_out/pytest_lev.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_lev.log
timeout 600 python tools/leverage_selfdistance.py > gpurun_out/levself.out 2>&1; echo "levself rc=$?"
timeout 120 python tools/probe_leverage.py 2>&1 | tail -3
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/lev_launches.csv python tools/probe_leverage.py > gpurun_out/ncu_lev.log 2>&1; echo "ncu rc=$?"
grep -E "leverage|eig" gpurun_out/lev_launches.csv | awk -F'","' '{print $5, $NF}' | tail -8
timeout 120 python tools/probe_eig.py 2>&1 | tail -6
