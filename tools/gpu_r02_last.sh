#!/bin/bash
# last evidence pass of round 2 (the prefill kernel changed): bench line, ncu launch list, full capture of the prefill kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu"
$CMD > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain_bench2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 3 -c 1 -f -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill.log 2>&1
echo "prefill capture rc=$?"
CVLLM_LIB= timeout 100 python tools/probe_prefill.py 2>&1 | tail -2
du -sh gpurun_out
