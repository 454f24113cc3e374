#!/bin/bash
# one GPU: ncu --set full of the fused GEMM + reduce kernel (world 1) -> profiles summary on the box
mkdir -p gpurun_out
timeout 120 python tools/gemm_reduce_once.py > gpurun_out/grs_once.log 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_reduce_kernel -o /tmp/grs python tools/gemm_reduce_once.py > gpurun_out/ncu_grs.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_grs.log
python tools/summarize_ncu_multi.py /tmp/grs.ncu-rep r02k gemm_reduce_kernel > gpurun_out/summ.log 2>&1; tail -3 gpurun_out/summ.log
cp profiles/r02k_gemm_reduce_kernel.json gpurun_out/ 2>/dev/null
ls -la /tmp/grs.ncu-rep
