#!/bin/bash
# ncu passes (launch list of the headline step, full capture of the prefill and decode kernels).
# Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu"
$CMD > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain_bench2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 3 -c 1 -f -o gpurun_out/prof_prefill $CMD > gpurun_out/ncu_prefill.log 2>&1
echo "prefill capture rc=$?"
DCMD="python tools/prof_decode.py"
$DCMD > gpurun_out/plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:decode_attn -s 4 -c 1 -f -o gpurun_out/prof_decode $DCMD > gpurun_out/ncu_decode.log 2>&1
echo "decode capture rc=$?"
tail -3 gpurun_out/plain_decode.log gpurun_out/ncu_prefill.log gpurun_out/ncu_decode.log
RCMD="python tools/probe_rope.py"
$RCMD > gpurun_out/plain_rope.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rope_qk -s 8 -c 1 -f -o gpurun_out/prof_rope $RCMD > gpurun_out/ncu_rope.log 2>&1
echo "rope capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:rope_qk -s 100 -c 1 -f -o gpurun_out/prof_rope_fused $RCMD > gpurun_out/ncu_rope_fused.log 2>&1
echo "fused rope capture rc=$?"
