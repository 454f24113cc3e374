#!/bin/bash
# tests + smoke + bench, then ncu launch list + full capture of the prefill kernel
bash tools/gpu_check.sh
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu"
$CMD > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain_bench2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 3 -c 1 -f -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill.log 2>&1
echo "prefill capture rc=$?"
