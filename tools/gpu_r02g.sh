#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scores.py tests/test_gpu_prepath.py tests/test_gpu_configs.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/leverage_stage_diff.py > gpurun_out/stage_diff.log 2>&1; tail -2 gpurun_out/stage_diff.log | cut -c1-400
echo "=== scoring stages"
timeout 300 python tools/prof_scoring.py 2>&1 | tail -5
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/scoring_launches.csv python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
grep -E "leverage_|chunk_scores" gpurun_out/scoring_launches.csv | awk -F'","' '{print substr($5,1,50), $NF}' | tail -7
echo "=== bench"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read()); c=d['components']
print({k: c.get(k) for k in ['compression_added_pct','compress_ms','compress_graph_ms','compress_graph_mtok_s','prefill_attn_kernel_ms','prefill_attn_kernel_alone_ms','compression_added_pct_4k','compression_added_pct_16k','leverage_from_k_ms','leverage_from_sketch_ms']})
print(d['value'], d['roofline']['frac'], d['e2e']['value'])
PY
tail -3 gpurun_out/bench.err
