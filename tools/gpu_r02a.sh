#!/bin/bash
# round-2 first GPU pass: tests + smoke + bench, sketch bit-parity ubenches, ncu captures of the scoring kernels
mkdir -p gpurun_out
bash tools/gpu_check.sh
echo "=== sketch ubench"
timeout 300 python tools/ubench/sketch_order.py > gpurun_out/sketch_order_py.log 2>&1; tail -12 gpurun_out/sketch_order_py.log
timeout 120 ./tools/ubench/sketch_order > gpurun_out/sketch_order.log 2>&1; cat gpurun_out/sketch_order.log
timeout 120 ./tools/ubench/sketch_umma > gpurun_out/sketch_umma.log 2>&1; cat gpurun_out/sketch_umma.log
echo "=== scoring kernels"
timeout 300 python tools/prof_scoring.py > gpurun_out/prof_scoring_plain.log 2>&1; cat gpurun_out/prof_scoring_plain.log | tail -5
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"chunk_|snap_|leverage_|sel_|store_all|compact_" -f -o gpurun_out/prof_scoring python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
echo "scoring capture rc=$?"; tail -3 gpurun_out/ncu_scoring.log
bash tools/gpu_launchlist.sh
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
