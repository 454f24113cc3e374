#!/bin/bash
# final evidence pass: tests, smoke, bench; ncu launch list + full captures (prefill, decode, rope, scoring kernels:
# summarised on the box, only the single-kernel reports travel); leverage self-distance against the reference
mkdir -p gpurun_out
bash tools/gpu_check.sh
timeout 900 python tools/leverage_selfdistance.py > gpurun_out/levself.out 2>&1; echo "levself rc=$?"
bash tools/gpu_profile.sh
timeout 300 python tools/prof_scoring.py --once > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:"chunk_|snap_|leverage_|sel_|store_all|compact_" -f -o /tmp/prof_scoring python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
echo "scoring capture rc=$?"
python tools/summarize_ncu_multi.py /tmp/prof_scoring.ncu-rep r02 scoring_kernels > gpurun_out/scoring_summary.txt 2>&1; cat gpurun_out/scoring_summary.txt
cp profiles/r02_scoring_kernels.json gpurun_out/
timeout 200 python tools/probe_eig.py > gpurun_out/probe_eig.txt 2>&1
du -sh gpurun_out
