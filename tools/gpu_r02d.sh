#!/bin/bash
# round-2 fourth GPU pass: full tests with the packed-softmax prefill + split scoring, chain timing, bench, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "=== scoring stages"
timeout 300 python tools/prof_scoring.py 2>&1 | tail -4
echo "=== bench"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
bash tools/gpu_launchlist.sh
du -sh gpurun_out
