#!/bin/bash
# round-2 second GPU pass: full tests, leverage self-distance + stage diff with the natural-k sketch, tcgen05 chunk
# scorer A/B, short-prompt overhead probe, e2e timeline, bench, scoring-kernel ncu summaries (made on the box: the raw
# reports are too large to travel)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
echo "=== leverage"
timeout 600 python tools/leverage_stage_diff.py > gpurun_out/stage_diff.log 2>&1; tail -3 gpurun_out/stage_diff.log
timeout 900 python tools/leverage_selfdistance.py > gpurun_out/levself.out 2>&1; echo "levself rc=$?"; tail -3 gpurun_out/levself.out
echo "=== overhead probe"
timeout 300 python tools/probe_overhead.py > gpurun_out/probe_overhead.log 2>&1; cat gpurun_out/probe_overhead.log | tail -40
echo "=== scoring stages (tcgen05 chunk scorer)"
timeout 300 python tools/prof_scoring.py > gpurun_out/prof_scoring_plain.log 2>&1; tail -4 gpurun_out/prof_scoring_plain.log
timeout 600 ncu --set full --clock-control none -k regex:"chunk_|snap_|leverage_|sel_|store_all|compact_" -f -o /tmp/prof_scoring python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
echo "scoring capture rc=$?"
python tools/summarize_ncu_multi.py /tmp/prof_scoring.ncu-rep r02 scoring_kernels > gpurun_out/scoring_summary.txt 2>&1; cat gpurun_out/scoring_summary.txt
cp profiles/r02_scoring_kernels.json gpurun_out/ 2>/dev/null
ls -la /tmp/prof_scoring.ncu-rep
echo "=== e2e timeline"
timeout 300 python tools/e2e_timeline.py > gpurun_out/e2e_timeline.log 2>&1; tail -30 gpurun_out/e2e_timeline.log
echo "=== bench"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
bash tools/gpu_launchlist.sh
du -sh gpurun_out
