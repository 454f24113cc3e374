#!/bin/bash
# multi-GPU pass (gpurun --gpus N): all-reduce parity + latency table (formal .sys barriers and the relaxed-flags
# variant for A/B), bench at N ranks (config-5 block inside), the 2-rank pytest
N=${1:-2}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
for n in ${2:-2} ${3:-} ${4:-}; do
  [ "$n" -le "$N" ] || continue
  timeout 600 bash -c "$(declare -f run); run $n 2951$n tools/tp_allreduce_check.py" > gpurun_out/tp_allreduce_sys_n$n.json 2> gpurun_out/tp_allreduce_sys_n$n.err
  echo "allreduce sys n=$n rc=$?"; python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/tp_allreduce_sys_n$n.json').read().strip().splitlines()[-1])
    print(' ok', d['ok'], {k: round(v, 1) for k, v in d['us'].items() if k.startswith('graph_') and isinstance(v, float)})
except Exception as ex:
    print(' parse failed', ex)
PY
  [ "$n" -eq "$N" ] && CVLLM_LIB=compactor-vllm_b200/lib/variants/lib_arrelaxed.so timeout 600 bash -c "$(declare -f run); run $n 2952$n tools/tp_allreduce_check.py" > gpurun_out/tp_allreduce_relaxed_n$n.json 2> gpurun_out/tp_allreduce_relaxed_n$n.err
  echo "allreduce relaxed n=$n rc=$?"
  timeout 900 bash -c "$(declare -f run); run $n 2953$n bench.py --gpus $n --steps 10 --warmup 3" > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n=$n rc=$?"; tail -c 3000 gpurun_out/bench_n$n.json; tail -3 gpurun_out/bench_n$n.err
done
timeout 600 python -m pytest tests/test_gpu_allreduce.py -m gpu -q > gpurun_out/pytest_allreduce.log 2>&1; echo "pytest allreduce rc=$?"; tail -3 gpurun_out/pytest_allreduce.log
du -sh gpurun_out
