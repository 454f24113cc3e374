#!/bin/bash
# round-2 third GPU pass: pipelined chunk scorer + split gather (tests, chain timing), eigensolver phase clocks,
# prefill f32x2 A/B, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scores.py tests/test_gpu_select.py tests/test_gpu_store.py tests/test_gpu_configs.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
echo "=== eig phases"
timeout 200 python tools/probe_eig.py 2>&1 | tail -6
echo "=== scoring stages"
timeout 300 python tools/prof_scoring.py 2>&1 | tail -4
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/scoring_launches.csv python tools/prof_scoring.py --once > gpurun_out/ncu_scoring.log 2>&1
python - <<'PY'
import csv, collections
rows = [l for l in open('gpurun_out/scoring_launches.csv') if not l.startswith('==')]
agg = collections.OrderedDict()
for x in csv.DictReader(rows):
    agg.setdefault(x['Kernel Name'].split('(')[0][:50], []).append(float(x['Metric Value'].replace(',', '')))
for k, v in agg.items():
    if 'cvllm' in k or 'kernel' in k: print(f"{k:52s} n={len(v):3d} last={v[-1]/1e3:8.1f} us")
PY
echo "=== prefill A/B (scalar vs f32x2 softmax math)"
for i in 1 2; do
timeout 200 python tools/probe_prefill.py 2>&1 | tail -2
CVLLM_LIB=compactor-vllm_b200/lib/variants/lib_f32x2.so timeout 200 python tools/probe_prefill.py 2>&1 | tail -2
done
echo "=== bench"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
du -sh gpurun_out
