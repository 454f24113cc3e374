#!/bin/bash
# fused o_proj GEMM + all-reduce at N ranks (gpurun --gpus N): kernel vs cuBLAS + NCCL, then the config-5 block of bench.py
N=${1:-2}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
for n in ${2:-$N} ${3:-} ${4:-}; do
  [ "$n" -le "$N" ] || continue
  timeout 300 bash -c "$(declare -f run); run $n 2961$n tools/gemm_reduce_bench.py" > gpurun_out/grs_n$n.json 2> gpurun_out/grs_n$n.err
  echo "gemm_reduce n=$n rc=$?"; grep "^{" gpurun_out/grs_n$n.json | tail -1; grep -i "error\|FAILED\|timeout rec" gpurun_out/grs_n$n.err | head -5
  timeout 600 bash -c "$(declare -f run); run $n 2962$n bench.py --gpus $n --steps 12 --warmup 3 --only-cfg5" > gpurun_out/cfg5_n$n.json 2> gpurun_out/cfg5_n$n.err
  echo "cfg5 n=$n rc=$?"; grep "^{" gpurun_out/cfg5_n$n.json | tail -1 | cut -c1-1800; tail -2 gpurun_out/cfg5_n$n.err
done
